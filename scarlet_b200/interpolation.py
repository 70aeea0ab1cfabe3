"""Fourier shifters and sinc interpolation used by the multi-resolution setup
(mirrors scarlet/interpolation.py:302-404; host, once per geometry)."""
import numpy as np


def mk_shifter(shape, real=False):
    """Unit phase ramps (exp(-2 pi i f)) of the two axes of an FFT of `shape`; raised to a power they shift by that
    many pixels (interpolation.py:302-340).  real=True: both axes sampled on rfft frequencies.  No cache: the
    reference's cache key ignores `real` (interpolation.py:319), so its two flavours could shadow each other."""
    fy = np.fft.rfftfreq(shape[0]) if real else np.fft.fftfreq(shape[0])
    fx = np.fft.rfftfreq(shape[1])
    return np.exp(-2j * np.pi * fy), np.exp(-2j * np.pi * fx)


def sinc_interp(images, coord_hr, coord_lr, angle=None, padding=3):
    """Samples of `images` (pixels at coord_lr) at the positions coord_hr by separable sinc interpolation
    (interpolation.py:342-372).  Like the reference, each image enters transposed (interpolation.py:370)."""
    if angle is not None:
        raise NotImplementedError("rotated frames are not supported")
    y_hr, x_hr = coord_hr
    y_lr, x_lr = coord_lr
    hy, hx = np.abs(y_lr[1] - y_lr[0]), np.abs(x_lr[1] - x_lr[0])
    assert hy != 0 and hx != 0
    wy = np.sinc((y_lr[None, :] - y_hr[:, None]) / hy)
    wx = np.sinc((x_lr[:, None] - x_hr[None, :]) / hx)
    return np.array([wy @ img.T @ wx for img in images])
