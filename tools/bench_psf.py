"""Throughput of the batched PSF matching (scl_match_psfs_f32, csrc/k_psf.cu): n PSF sets x C bands, 41x41 stamps,
difference kernels + 128-grid spectra, host buffers in and out; next to the host container path
(scarlet_b200.fft.match_psfs + _native.kernel_spectrum, what Observation.match did before) on a sample.

    python tools/bench_psf.py --sets 2000 --bands 5
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def moffat(shape, alpha, beta):
    y, x = np.indices(shape, dtype=np.float64)
    r2 = (y - shape[0] // 2) ** 2 + (x - shape[1] // 2) ** 2
    p = (1 + r2 / alpha ** 2) ** -beta
    return p / p.sum()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sets", type=int, default=2000)
    ap.add_argument("--bands", type=int, default=5)
    ap.add_argument("--stamp", type=int, default=41)
    ap.add_argument("--frame", type=int, default=64)
    ap.add_argument("--host-sample", type=int, default=50)
    args = ap.parse_args()
    from scarlet_b200 import _native, fft, psf as spsf
    _native.require_device()
    rng = np.random.default_rng(0)
    k = args.stamp
    base = np.stack([moffat((k, k), a, 3.0) for a in np.linspace(2.0, 3.5, 16)]).astype(np.float32)
    obs = base[rng.integers(0, 16, (args.sets, args.bands))]
    model = spsf.generate_psf_image(spsf.gaussian, (k, k), sigma=.9).image[None].astype(np.float32)
    _native.match_psfs(obs[:8], model, frame_shape=(args.frame, args.frame))            # warm-up
    times = []
    for _ in range(3):
        t0 = time.perf_counter()
        diff, spec = _native.match_psfs(obs, model, frame_shape=(args.frame, args.frame))
        times.append(time.perf_counter() - t0)
    t_dev = min(times)
    ns = min(args.host_sample, args.sets)
    t0 = time.perf_counter()
    worst = 0.0
    for i in range(ns):
        d = fft.match_psfs(fft.Fourier(obs[i]), fft.Fourier(model)).image
        s = _native.kernel_spectrum(d, args.frame, args.frame)
        worst = max(worst, float(np.abs(d - diff[i]).max() / np.abs(d).max()))
    t_host = (time.perf_counter() - t0) / ns
    print(json.dumps({"op": "match_psfs + kernel spectra", "sets": args.sets, "bands": args.bands, "stamp": k,
                      "device_call_s": t_dev, "sets_per_s": args.sets / t_dev,
                      "bytes_out": int(diff.nbytes + spec.nbytes),
                      "host_numpy_s_per_set": t_host, "host_sets_per_s": 1.0 / t_host,
                      "max_rel_diff_vs_host_f32_fft": worst}))


if __name__ == "__main__":
    main()
