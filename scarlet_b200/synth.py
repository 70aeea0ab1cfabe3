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
