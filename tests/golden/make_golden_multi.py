"""Generate tests/golden/ref_golden_multi.npz: the reference package (/root/reference, imported unmodified through
ref_shims.py) on config 1 (data/hsc_cosmos.npz) with the first two catalogue rows as MultiComponentSource
(source.py:537-641, flux_percentiles default [25]) and the rest as ExtendedSource.  Run in the build container:
  python tests/golden/make_golden_multi.py
"""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_shims  # noqa: E402

scarlet = ref_shims.install()
from scarlet.cache import Cache  # noqa: E402

N_MULTI, MAX_ITER = 2, 40
Cache._cache.clear()
d = np.load("/root/reference/data/hsc_cosmos.npz")
images, psfs, cat = d["images"], d["psfs"], d["catalog"]
mpsf = scarlet.psf.generate_psf_image(scarlet.psf.gaussian, psfs[0].shape, sigma=.9).image
frame = scarlet.Frame(images.shape, psfs=mpsf[None])
obs = scarlet.Observation(images, psfs=psfs).match(frame)
bg_rms = np.ones(len(images), dtype=images.dtype) * 0.1
srcs = []
for i, row in enumerate(cat):
    cls = scarlet.MultiComponentSource if i < N_MULTI else scarlet.ExtendedSource
    srcs.append(cls(frame, (row["y"], row["x"]), obs, bg_rms))
blend = scarlet.Blend(srcs, obs)
out = dict(n_multi=N_MULTI, max_iter=MAX_ITER)
out["n_comp"] = np.array([s.n_components if hasattr(s, "n_components") else 1 for s in srcs])
out["init_seds"] = np.array([c.sed for c in blend.components])
out["init_morphs"] = np.array([c.morph for c in blend.components])
out["init_centers"] = np.array([s.pixel_center for s in srcs])
blend.fit(MAX_ITER)
out["mse"] = np.array(blend.mse, dtype=np.float64)
out["seds"] = np.array([c.sed for c in blend.components])
out["morphs"] = np.array([c.morph for c in blend.components])
out["centers"] = np.array([s.pixel_center for s in srcs])
out["model"] = blend.get_model()
print("components", len(blend.components), "iterations", blend.it, "converged", blend.converged,
      "mse", blend.mse[0], blend.mse[-1])
np.savez_compressed(os.path.join(HERE, "ref_golden_multi.npz"), **out)
