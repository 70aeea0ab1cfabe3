"""Reproducer: what the reference's ROTATED LowResObservation branch (observation.py:443-454,483-503, sinc_shift
observation.py:345-403 with axes=(1, 2)) does when run UNMODIFIED from /root/reference through ref_shims.py on a 10-degree
rotated low-resolution frame that lies entirely inside the model frame (rotated variant).
Run in the build container:   python tests/golden/repro_rotated_lowres.py
Prints one line per attempt (cold operator cache / cache primed by a real=True shifter of the same shape) and writes
them to tests/golden/repro_rotated_lowres.txt."""
import os
import sys
import traceback
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_shims  # noqa: E402
from wcs_stub import make_pair  # noqa: E402

scarlet = ref_shims.install()
import autograd.numpy as agnp  # noqa: E402
from scarlet.cache import Cache  # noqa: E402
from scarlet import interpolation  # noqa: E402

_arr = np.array


def ragged_array(obj, *a, **k):
    try:
        return _arr(obj, *a, **k)
    except ValueError:
        out = np.empty(len(obj), dtype=object)
        for i, o in enumerate(obj):
            out[i] = o
        return out


agnp.array = ragged_array
psf = scarlet.psf
CH = ["h0", "h1", "h2", "l0", "l1", "l2", "l3", "l4"]


def attempt(prime_cache, angle=10.0, hr=(64, 64), lr=(8, 8)):
    Cache._cache.clear()
    whr, wlr = make_pair(hr, lr, scale=4.0, angle_deg=angle, lr_offset=(0.3, -0.2))
    mpsf = psf.generate_psf_image(psf.gaussian, (15, 15), sigma=.9).image.astype(np.float32)
    frame = scarlet.Frame((8,) + hr, wcs=whr, psfs=mpsf[None], channels=CH)
    lr_psfs = np.array([psf.generate_psf_image(psf.gaussian, (9, 9), sigma=.8 + .05 * b).image
                        for b in range(5)]).astype(np.float32)
    olr = scarlet.LowResObservation(np.zeros((5,) + lr, np.float32), wcs=wlr, psfs=lr_psfs, channels=CH[3:])
    stage = "match"
    try:
        if prime_cache:
            # the same shape asked for with real=True first: mk_shifter caches by shape only (interpolation.py:319-338)
            import scarlet.fft as sfft
            shape = sfft._get_fft_shape(frame.psfs, np.zeros(frame.shape), padding=3, axes=[-2, -1], max=True)
            interpolation.mk_shifter(shape, real=True)
        olr.match(frame)
        stage = "_render"
        model = np.random.default_rng(0).random((8,) + hr).astype(np.float32)
        out = olr._render(model)
        return "rotated %.0f deg, cache %s: match + _render ran, isrot=%s, output shape %s, finite=%s" % (
            angle, "primed" if prime_cache else "cold", olr.isrot, out.shape, bool(np.isfinite(out).all()))
    except Exception as e:      # noqa: BLE001
        tb = traceback.extract_tb(e.__traceback__)[-1]
        return "rotated %.0f deg, cache %s: %s in %s -> %s: %s (%s:%d)" % (
            angle, "primed" if prime_cache else "cold", type(e).__name__, stage, tb.name, str(e)[:160],
            os.path.basename(tb.filename), tb.lineno)


if __name__ == "__main__":
    lines = [attempt(False), attempt(True), attempt(False, lr=(10, 10)), attempt(True, lr=(10, 10))]
    for ln in lines:
        print(ln)
    with open(os.path.join(HERE, "repro_rotated_lowres.txt"), "w") as f:
        f.write("\n".join(lines) + "\n")
