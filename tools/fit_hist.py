"""Convergence profile of the config-3 workload: iteration histogram, device time vs summed kernel time.
  python tools/fit_hist.py [B] [check_every]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from scarlet_b200 import synth            # noqa: E402
from scarlet_b200.batch import BlendBatch  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
ce = int(sys.argv[2]) if len(sys.argv) > 2 else 8
d = synth.make_blends(range(B), device="cpu")
bb = BlendBatch(d["images"], d["psfs"], d["model_psf"], d["centers"], d["bg_rms"], weights=d["weights"])
bb.snapshot()
for prof in (False, True):
    bb.restore()
    info = bb.fit_device(200, 1e-2, check_every=ce, profile=prof)
    print("profile=%s device_ms=%.2f steps=%d launches=%d kernel_ms=%s sum=%.2f" % (
        prof, info["device_ms"], info["steps"], info["launches"], np.round(info["kernel_ms"], 2), sum(info["kernel_ms"])))
n = bb.download(morph=False)["n_iter"]
print("n_iter: mean %.1f median %d p90 %d p99 %d max %d; blends at 200: %d" % (
    n.mean(), np.median(n), np.percentile(n, 90), np.percentile(n, 99), n.max(), (n >= 200).sum()))
h = np.bincount(np.minimum(n, 200) // 10)
print("hist by 10:", h.tolist())
act = [(n > s).sum() for s in range(0, 200, 8)]
print("active at step 0,8,..:", act)
