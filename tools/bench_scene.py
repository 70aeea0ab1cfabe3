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
from scarlet_b200 import synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--sources", type=int, default=200)
    ap.add_argument("--bands", type=int, default=6)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--approx", action="store_true")
    ap.add_argument("--debug-flags", type=int, default=0, help="scl_batch_debug_flags of the batch (experiments)")
    ap.add_argument("--drop", default="", help="experiment: clear a constraint flag on every component (symmetric, monotonic, kspace)")
    a = ap.parse_args()
    C, Ny, Nx, ns = a.bands, a.size, a.size, a.sources
    t0 = time.time()
    nat.require_device()
    batch, _ = synth.scene_arrays(a.size, a.sources, a.bands)
    K = len(batch["sed"])
    if a.drop:
        bit = {"symmetric": nat.F_SYMMETRIC, "monotonic": nat.F_MONOTONIC, "kspace": nat.F_KSPACE}[a.drop]
        batch["cflags"] = [f & ~bit for f in batch["cflags"]]
    t_gen = time.time() - t0
    t0 = time.time()
    bt = nat.Batch(**batch)
    if a.debug_flags:
        bt.debug_flags(a.debug_flags)
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
    # slots 5..7: longest symmetry phase / sweep / CTA of a launch (max over the three launches), in ms at 1.9 GHz
    longest = dict(zip(["symmetry_ms", "sweep_ms", "cta_ms"], [round(float(c) / 1.9e6, 3) for c in clk[5:8]]))
    bt.close()
    N = Ny * Nx
    alg = 4 * N * (2 * K + 2 * C) + 4 * C * 41 * 41 + 8 * K * C         # DESIGN.md B_iter
    ms = info["device_ms"] / info["steps"]
    km = [k / info["steps"] for k in info["kernel_ms"]]
    print(json.dumps(dict(workload="config4-like scene", C=C, Ny=Ny, Nx=Nx, sources=ns, components=K,
                          approximate_L=a.approx, iters=info["steps"], ms_per_iter=ms,
                          kernel_ms=dict(render_conv_adjoint=km[0], lipschitz=km[1], constraints=km[2], converge=km[3]),
                          constraint_phase_share=phases, constraint_longest=longest, alg_bytes_per_iter=alg, gbs=alg / ms / 1e6, launches=info["launches"],
                          logL=[float(v) for v in out["logL"][0, :info["steps"]]],
                          gen_s=t_gen, tables_upload_s=t_up)))


if __name__ == "__main__":
    main()
