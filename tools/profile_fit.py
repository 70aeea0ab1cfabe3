"""Short fixed-iteration fit for profilers: B synthetic config-3 blends, `iters` iterations with every blend active.
  python tools/profile_fit.py [B] [iters]
ncu: skip the construction-time launches with -s and capture one iteration's kernels (see profiles/README.md)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from scarlet_b200 import synth            # noqa: E402
from scarlet_b200.batch import BlendBatch  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 12
d = synth.make_blends(range(B), device="cpu")
bb = BlendBatch(d["images"], d["psfs"], d["model_psf"], d["centers"], d["bg_rms"], weights=d["weights"])
if os.environ.get("SCL_PHASES"):
    bb._dev.phase_clocks(True)
t0 = time.perf_counter()
info = bb.fit_device(iters, 0.0, check_every=1000, profile=True)
if os.environ.get("SCL_PHASES"):
    clk = bb._dev.phase_clocks(False)
    tot = sum(clk) or 1
    print("K3 phase clocks per CTA: " + ", ".join("%s %.0f (%.0f%%)" % (n, c / (bb.K * info["steps"]), 100 * c / tot) for n, c in
          zip(["step", "centre+symmetry", "list", "sweep", "finish", "x5", "x6", "x7"], clk)))
print("B=%d K=%d iters=%d device_ms=%.3f kernel_ms/launch=%s" % (
    bb.B, bb.K, info["steps"], info["device_ms"], [round(v / info["steps"], 4) for v in info["kernel_ms"]]))
