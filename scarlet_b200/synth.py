"""Synthetic blends of BASELINE config 3 (SURVEY 8(d) "Synthetic inputs"): C-band Ny x Nx scenes with
2-8 elliptical Sersic-like sources, per-band Gaussian PSFs, unit Gaussian noise.

This is input generation, not the hot path.  The per-blend random draws follow one fixed order from
`numpy.random.default_rng(20260411 + blend_index)` (the same order tests/helpers.synthetic_blend uses, so
both describe the same scenes); the PSF convolution of the true scene is a float64 FFT product done with
torch on whatever device is available.
"""
import numpy as np

from .psf import generate_psf_image, gaussian

SEED0 = 20260411


def config3_psfs(C=5, psf_size=41, dtype=np.float32):
    """(model_psf [1,k,k], psfs [C,k,k]): sigma .9 model PSF and sigma_b = 1.2 + .15 b band PSFs, normalised."""
    mpsf = generate_psf_image(gaussian, (psf_size, psf_size), sigma=.9).image
    psfs = np.array([generate_psf_image(gaussian, (psf_size, psf_size), sigma=1.2 + .15 * b).image for b in range(C)])
    return mpsf[None].astype(dtype), psfs.astype(dtype)


def _psfs64(C, psf_size):
    return np.array([generate_psf_image(gaussian, (psf_size, psf_size), sigma=1.2 + .15 * b).image for b in range(C)])


def draw_scene(index, C=5, Ny=64, Nx=64, kmin=2, kmax=8):
    """Random parameters of blend `index`: dict(centers [K,2], re, n, q, pa [K], sed [K,C]) and the rng,
    positioned right before the noise draw."""
    rng = np.random.default_rng(SEED0 + int(index))
    K = int(rng.integers(kmin, kmax + 1))
    centers = []
    while len(centers) < K:
        c = rng.uniform(12, min(Ny, Nx) - 12, 2)
        if all(np.hypot(*(c - o)) >= 6 for o in centers):
            centers.append(c)
    re, n, q, pa, sed = [], [], [], [], []
    for _ in centers:
        re.append(rng.uniform(1.5, 5)); n.append(rng.uniform(.7, 2.5))
        q.append(rng.uniform(.5, 1)); pa.append(rng.uniform(0, np.pi))
        sed.append(rng.dirichlet(np.ones(C)) * 10 ** rng.uniform(np.log10(50), np.log10(5000)))
    return dict(centers=np.array(centers), re=np.array(re), n=np.array(n), q=np.array(q), pa=np.array(pa),
                sed=np.array(sed)), rng


def make_blends(indices, C=5, Ny=64, Nx=64, kmin=2, kmax=8, psf_size=41, device=None, chunk=512):
    """Images and catalogues for the blends `indices`.

    Returns dict(images f32 [B,C,Ny,Nx], weights f32 [B,C,Ny,Nx] (= 1/sigma_n^2 = 1), bg_rms f32 [C],
    model_psf f32 [1,k,k], psfs f32 [C,k,k], centers: list of int [K_b,2] pixel catalogues (y, x))."""
    import torch
    if device is None:
        device = "cuda" if torch.cuda.is_available() else "cpu"
    indices = list(indices)
    B = len(indices)
    mpsf, psfs = config3_psfs(C, psf_size)
    # PSF spectra on a periodic grid with the centre tap at index 0 ('same' convolution)
    F = 1
    while F < max(Ny, Nx) + psf_size:
        F *= 2
    g = np.zeros((C, F, F))
    idx = (np.arange(psf_size) - psf_size // 2) % F
    g[:, idx[:, None], idx[None, :]] = _psfs64(C, psf_size)
    P = torch.fft.rfft2(torch.from_numpy(g).to(device))
    yy = torch.arange(Ny, dtype=torch.float64, device=device)[:, None]
    xx = torch.arange(Nx, dtype=torch.float64, device=device)[None, :]
    images = np.empty((B, C, Ny, Nx), np.float32)
    catalog = []
    for s in range(0, B, chunk):
        e = min(B, s + chunk)
        par = np.zeros((e - s, kmax, 6))
        sed = np.zeros((e - s, kmax, C))
        noise = np.empty((e - s, C, Ny, Nx))
        for i, ind in enumerate(indices[s:e]):
            sc, rng = draw_scene(ind, C, Ny, Nx, kmin, kmax)
            K = len(sc["re"])
            par[i, :K] = np.stack([sc["centers"][:, 0], sc["centers"][:, 1], sc["re"], sc["n"], sc["q"], sc["pa"]], 1)
            par[i, K:, 2:5] = 1.0
            sed[i, :K] = sc["sed"]
            noise[i] = rng.normal(0, 1, (C, Ny, Nx))
            catalog.append(np.array([[int(round(c[0])), int(round(c[1]))] for c in sc["centers"]], dtype=np.int32))
        p = torch.from_numpy(par).to(device)[..., None, None]         # [b,k,6,1,1]
        dy, dx = yy - p[:, :, 0], xx - p[:, :, 1]
        cpa, spa = torch.cos(p[:, :, 5]), torch.sin(p[:, :, 5])
        u = dx * cpa + dy * spa
        v = (-dx * spa + dy * cpa) / p[:, :, 4]
        morph = torch.exp(-(torch.hypot(u, v) / p[:, :, 2]) ** (1.0 / p[:, :, 3]))      # [b,k,Ny,Nx]
        scene = torch.einsum("bkc,bkyx->bcyx", torch.from_numpy(sed).to(device), morph)
        conv = torch.fft.irfft2(torch.fft.rfft2(scene, s=(F, F)) * P, s=(F, F))[..., :Ny, :Nx]
        images[s:e] = (conv.cpu().numpy() + noise).astype(np.float32)
    return dict(images=images, weights=np.ones_like(images), bg_rms=np.ones(C, np.float32),
                model_psf=mpsf, psfs=psfs, centers=catalog)


def scene_arrays(size=512, sources=200, bands=6, psf_size=41, seed=SEED0, multi_fraction=.7):
    """One synthetic LSST-like scene of BASELINE config 4: `bands` x size x size, `sources` sources with
    centres at least 10 px apart, 70 % as two-component sources (a core above half the peak + the rest, sharing one
    centre like MultiComponentSource), 30 % single-component, monotonic + symmetric constraints on all of them.
    Components start from the generator's own (thresholded) profiles: a timing workload for the iteration; the
    source-initialisation path is not part of it.  Returns the arrays `_native.Batch` takes (one blend) plus
    `multi` [K] (component belongs to a two-component source), `psfs`, `mpsf`, `diff` (difference kernels)."""
    from . import _native as nat
    from . import fft as sfft
    from scipy.signal import fftconvolve
    C, Ny, Nx = bands, size, size
    rng = np.random.default_rng(seed)
    mpsf, psfs = config3_psfs(C, psf_size)
    diff = sfft.match_psfs(sfft.Fourier(psfs), sfft.Fourier(mpsf)).image.astype(np.float32)
    centers = []
    while len(centers) < sources:
        c = rng.uniform(12, size - 12, 2)
        if all(np.hypot(*(c - o)) >= 10 for o in centers):
            centers.append(c)
    yy, xx = np.indices((Ny, Nx), dtype=np.float32)
    sed, morph, cflags, center, src_first, multi = [], [], [], [], [], []
    truth = np.zeros((C, Ny, Nx), np.float32)
    base = nat.F_CONSTRAINED | nat.F_SYMMETRIC | nat.F_MONOTONIC
    for c in centers:
        re, n, q, pa = rng.uniform(1.5, 5), rng.uniform(.7, 2.5), rng.uniform(.5, 1), rng.uniform(0, np.pi)
        dy, dx = yy - c[0], xx - c[1]
        u = dx * np.cos(pa) + dy * np.sin(pa)
        v = (-dx * np.sin(pa) + dy * np.cos(pa)) / q
        m = np.exp(-(np.hypot(u, v) / re) ** (1 / n)).astype(np.float32)
        m[m < 1e-3] = 0
        s = (rng.dirichlet(np.ones(C)) * 10 ** rng.uniform(np.log10(50), np.log10(5000))).astype(np.float32)
        truth += s[:, None, None] * m[None]
        pc = (int(round(c[0])), int(round(c[1])))
        first = len(sed)
        if rng.random() < multi_fraction:
            core = np.where(m >= .5 * m.max(), m, 0).astype(np.float32)
            for part, frac in ((m - .5 * core, .6), (core, .4)):
                sed.append(s * frac); morph.append(part / part.max()); center.append(pc)
                cflags.append(base | nat.F_SHARED_CENTER); src_first.append(first); multi.append(True)
        else:
            sed.append(s); morph.append(m / m.max()); center.append(pc)
            cflags.append(base | nat.F_KSPACE); src_first.append(first); multi.append(False)
    images = np.array([fftconvolve(truth[b], psfs[b], mode="same") for b in range(C)], dtype=np.float32)
    images += rng.normal(0, 1, images.shape).astype(np.float32)
    K = len(sed)
    batch = dict(C=C, Ny=Ny, Nx=Nx, comp_off=[0, K], sed=np.array(sed), morph=np.array(morph), cflags=cflags,
                 center=center, shift=np.zeros((K, 2)), conv=3 * np.ones(K), images=images[None],
                 weights=np.ones((1, C, Ny, Nx), np.float32), kernel_spec=nat.kernel_spectrum(diff, Ny, Nx),
                 ker_idx=np.arange(C), cw=mpsf, cw_idx=[0], src_first=src_first,
                 ker_half=nat.kernel_half(psf_size, psf_size, Ny, Nx))
    return batch, dict(multi=np.array(multi), psfs=psfs, mpsf=mpsf, diff=diff, images=images)
