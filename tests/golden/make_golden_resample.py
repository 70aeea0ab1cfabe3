"""Generate tests/golden/ref_golden_resample.npz: the reference package (/root/reference, imported unmodified through
ref_shims.py) on interpolation.fft_resample (interpolation.py:408-448) and update.translation (update.py:159-167)
with the lanczos, bilinear, cubic-spline family and quintic kernels, plus update.threshold / measurement.threshold
(update.py:85-103, measurement.py:96-113) on a blob over a noise floor.  Run in the build container:
  python tests/golden/make_golden_resample.py
"""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_shims  # noqa: E402

scarlet = ref_shims.install()
from scarlet import interpolation, update, measurement  # noqa: E402

rng = np.random.default_rng(42)
out = {}
img = np.zeros((23, 31))
img[4:17, 6:25] = rng.random((13, 19))
y, x = np.indices(img.shape)
img += 3 * np.exp(-((y - 10.3) ** 2 + (x - 14.8) ** 2) / 6.0)
out["img"] = img
shifts = np.array([(.217, -.026), (-.691, .321), (.5, .5), (-.999, .001), (0., 0.), (.75, -.4)])
out["shifts"] = shifts
kernels = ["lanczos", "bilinear", "catmull_rom", "mitchel_netravali", "cubic_spline", "quintic_spline"]
for name in kernels:
    fn = getattr(interpolation, name)
    out["resample_" + name] = np.array([interpolation.fft_resample(img, dy, dx, kernel=fn) for dy, dx in shifts])
    out["taps_" + name] = np.array([np.concatenate(fn(d)) for d in (.217, -.691, .5, 0.)])


class Comp:
    pass


res = []
for direction in (1, -1):
    c = Comp()
    c.morph = img.astype(np.float32)
    c.shift = (.31, -.44)
    update.translation(c, direction)
    res.append(c.morph)
out["translation"] = np.array(res)

rs = np.random.RandomState(0)
noise = rs.rand(21, 21) * 2
signal = np.zeros(noise.shape)
signal[7:14, 7:14] = scarlet.psf.generate_psf_image(scarlet.psf.gaussian, (21, 21), normalize=False, amplitude=10,
                                                    sigma=3)[7:14, 7:14].image
c = Comp()
c.morph = (signal + noise).astype(np.float32)
out["thresh_in"] = c.morph.copy()
t, b = measurement.threshold(c.morph)
out["thresh_value"], out["thresh_bins"] = np.float64(t), b
update.threshold(c)
out["thresh_out"] = c.morph
bb = c.bboxes["thresh"]
out["thresh_box"] = np.array([bb.bottom, bb.top, bb.left, bb.right])
np.savez_compressed(os.path.join(HERE, "ref_golden_resample.npz"), **out)
print({k: np.shape(v) for k, v in out.items()})
