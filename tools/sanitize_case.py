"""One small process that drives every hot kernel once, for compute-sanitizer (SURVEY 5 "race detection"):
a 48-blend config-3 batch (K1 scl_k_forward_adjoint, K2, K3 scl_k_prox with the streamed sweep and its cp.async ring,
scl_k_mono_plan, scl_k_init_extended) for 3 iterations, and a 96x96 scene with 10 sources forced through scene mode
(scl_k_scene_*, scl_k_scene_bbox / gram_boxes, scl_k_prox_wide, scl_k_source_center_wide) for 2 iterations.
  compute-sanitizer --tool racecheck python tools/sanitize_case.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from scarlet_b200 import synth, _native as nat  # noqa: E402
from scarlet_b200.batch import BlendBatch        # noqa: E402

d = synth.make_blends(range(48), device="cpu")
bb = BlendBatch(d["images"], d["psfs"], d["model_psf"], d["centers"], d["bg_rms"], weights=d["weights"] * 1.5)
info = bb.fit_device(3, 0.0, check_every=1000)
print("config 3: %d blends, %d components, %d launches" % (bb.B, bb.K, info["launches"]))
bb.close()
batch, _ = synth.scene_arrays(96, 10, 6)
bt = nat.Batch(**batch)
info = bt.fit(2, 0.0)
print("scene: %d components, %d launches" % (len(batch["sed"]), info["launches"]))
bt.close()
