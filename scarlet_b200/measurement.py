"""Centre measurement (mirrors scarlet/measurement.py:3-94), evaluated on the device."""
import numpy as np

from . import _native
from . import schedule


def max_pixel(morph, center=None, window=None):
    """Pixel with the maximum flux in the 5x5 window around `center` (measurement.py:3-29)."""
    if window is not None:
        raise NotImplementedError("max_pixel: only the default 5x5 window runs on the device")
    if schedule.recording() is not None:
        if center is None:
            raise NotImplementedError("inside the CUDA loop max_pixel starts from the component's pixel_center")
        schedule.record("max_pixel")
        return schedule.LiveCenter(center)
    if center is None:
        center = (morph.shape[0] // 2, morph.shape[1] // 2)
    c, _ = _native.measure(np.asarray(morph, dtype=np.float32)[None], [(int(center[0]), int(center[1]))],
                           None, True, False)
    return (int(c[0, 0]), int(c[0, 1]))


def psf_weighted_centroid(morph, psf, pixel_center):
    """PSF-weighted first moments around `pixel_center` (measurement.py:32-94).
    Returns (rounded centre, (dy, dx))."""
    if schedule.recording() is not None:
        schedule.record("centroid")          # the device weights with the frame's centroid weight (source.py:389-397)
        return schedule.LiveCenter(pixel_center), schedule.LiveShift((0., 0.))
    c, s = _native.measure(np.asarray(morph, dtype=np.float32)[None],
                           [(int(pixel_center[0]), int(pixel_center[1]))], psf, False, True)
    return (int(c[0, 0]), int(c[0, 1])), (float(s[0, 0]), float(s[0, 1]))


def threshold(morph):
    """Cut-off below which a morphology is fitting noise: 10^(left edge of the last empty bin of the log10 histogram
    of its positive pixels), and the number of bins used (measurement.py:96-113).  On the device (csrc/k_update.cu)."""
    th, bins = _native.threshold(np.ascontiguousarray(morph, dtype=np.float32)[None])
    return (0 if th[0] == 0 else float(th[0])), int(bins[0])
