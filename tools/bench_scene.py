"""BASELINE config 4 shape class on the device: one synthetic LSST-like 6-band 512x512 scene with 200 sources
(70 % two-component, 30 % single-component => K ~ 340), monotonic + symmetric constraints, 41x41 PSFs
(SURVEY.md 8(d)).  Device time per iteration and per kernel group of scene mode (k_scene.cu), exact or approximate
Lipschitz constants.  Components are initialised from the generator's own profiles (a timing workload: the
source-initialisation path is not part of the iteration).

  python tools/bench_scene.py [--size 512] [--sources 200] [--iters 10] [--approx]
"""
import argparse
import json
import os
import sys
import time
import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scarlet_b200 import _native as nat  # noqa: E402
from scarlet_b200 import psf as spsf  # noqa: E402
from scarlet_b200 import fft as sfft  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--sources", type=int, default=200)
    ap.add_argument("--bands", type=int, default=6)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--approx", action="store_true")
    a = ap.parse_args()
    C, Ny, Nx, ns = a.bands, a.size, a.size, a.sources
    rng = np.random.default_rng(20260411)
    t0 = time.time()
    nat.require_device()
    mpsf = spsf.generate_psf_image(spsf.gaussian, (41, 41), sigma=.9).image.astype(np.float32)
    psfs = np.array([spsf.generate_psf_image(spsf.gaussian, (41, 41), sigma=1.2 + .15 * b).image for b in range(C)],
                    dtype=np.float32)
    diff = sfft.match_psfs(sfft.Fourier(psfs), sfft.Fourier(mpsf[None])).image.astype(np.float32)
    spec = nat.kernel_spectrum(diff, Ny, Nx)
    centers = []
    while len(centers) < ns:
        c = rng.uniform(12, Ny - 12, 2)
        if all(np.hypot(*(c - o)) >= 10 for o in centers):
            centers.append(c)
    yy, xx = np.indices((Ny, Nx), dtype=np.float32)
    sed, morph, cflags, center, src_first = [], [], [], [], []
    images = np.zeros((C, Ny, Nx), np.float32)
    from scipy.signal import fftconvolve
    truth = np.zeros((C, Ny, Nx), np.float32)
    for i, c in enumerate(centers):
        re, n, q, pa = rng.uniform(1.5, 5), rng.uniform(.7, 2.5), rng.uniform(.5, 1), rng.uniform(0, np.pi)
        dy, dx = yy - c[0], xx - c[1]
        u = dx * np.cos(pa) + dy * np.sin(pa)
        v = (-dx * np.sin(pa) + dy * np.cos(pa)) / q
        m = np.exp(-(np.hypot(u, v) / re) ** (1 / n)).astype(np.float32)
        m[m < 1e-3] = 0
        s = (rng.dirichlet(np.ones(C)) * 10 ** rng.uniform(np.log10(50), np.log10(5000))).astype(np.float32)
        truth += s[:, None, None] * m[None]
        pc = (int(round(c[0])), int(round(c[1])))
        multi = rng.random() < .7
        first = len(sed)
        base = nat.F_CONSTRAINED | nat.F_SYMMETRIC | nat.F_MONOTONIC
        if multi:
            core = np.where(m >= .5 * m.max(), m, 0).astype(np.float32)
            for part, frac in ((m - .5 * core, .6), (core, .4)):
                sed.append(s * frac); morph.append(part / part.max()); center.append(pc)
                cflags.append(base | nat.F_SHARED_CENTER); src_first.append(first)
        else:
            sed.append(s); morph.append(m / m.max()); center.append(pc)
            cflags.append(base | nat.F_KSPACE); src_first.append(first)
    for b in range(C):
        images[b] = fftconvolve(truth[b], psfs[b], mode="same")
    images += rng.normal(0, 1, images.shape).astype(np.float32)
    K = len(sed)
    t_gen = time.time() - t0
    t0 = time.time()
    bt = nat.Batch(C, Ny, Nx, comp_off=[0, K], sed=np.array(sed), morph=np.array(morph), cflags=cflags, center=center,
                   shift=np.zeros((K, 2)), conv=np.zeros(K), images=images[None],
                   weights=np.ones((1, C, Ny, Nx), np.float32), kernel_spec=spec, ker_idx=np.arange(C),
                   cw=mpsf[None], cw_idx=[0], src_first=src_first, ker_half=nat.kernel_half(41, 41, Ny, Nx))
    t_up = time.time() - t0
    bt.fit(2, 0.0, approximate_L=a.approx)                       # warm-up
    info = bt.fit(a.iters, 0.0, approximate_L=a.approx, profile=True)
    out = bt.download()
    # share of the constraint kernel's phases (clock64 sums over CTAs)
    bt.phase_clocks(True)
    bt.fit(3, 0.0, approximate_L=a.approx)
    clk = bt.phase_clocks(False)
    tot = float(sum(clk[:5])) or 1.0
    phases = dict(zip(["gradient_step", "centre_symmetry", "level_list", "monotonic", "normalise_writeback"],
                      [round(c / tot, 3) for c in clk[:5]]))
    bt.close()
    N = Ny * Nx
    alg = 4 * N * (2 * K + 2 * C) + 4 * C * 41 * 41 + 8 * K * C         # DESIGN.md B_iter
    ms = info["device_ms"] / info["steps"]
    km = [k / info["steps"] for k in info["kernel_ms"]]
    print(json.dumps(dict(workload="config4-like scene", C=C, Ny=Ny, Nx=Nx, sources=ns, components=K,
                          approximate_L=a.approx, iters=info["steps"], ms_per_iter=ms,
                          kernel_ms=dict(render_conv_adjoint=km[0], lipschitz=km[1], constraints=km[2], converge=km[3]),
                          constraint_phase_share=phases, alg_bytes_per_iter=alg, gbs=alg / ms / 1e6, launches=info["launches"],
                          logL=[float(v) for v in out["logL"][0, :info["steps"]]],
                          gen_s=t_gen, tables_upload_s=t_up)))


if __name__ == "__main__":
    main()
