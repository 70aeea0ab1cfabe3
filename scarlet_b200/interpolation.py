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
    y_hr, x_hr = coord_hr
    y_lr, x_lr = coord_lr
    hy, hx = np.abs(y_lr[1] - y_lr[0]), np.abs(x_lr[1] - x_lr[0])
    assert hy != 0 and hx != 0
    if angle is not None:
        # Rotated grids (interpolation.py:374-404), angle = (cos, sin).  For every high-resolution row a the images are
        # moved by (-y_a cos / hy, -y_a sin / hx) pixels with a Fourier phase ramp on the zero-padded box, then
        # sample (a, b) is the double sinc sum over the moved image at (-x_b sin, x_b cos).
        from . import fft as sfft
        cos, sin = angle
        images = np.asarray(images)
        box = sfft._get_fft_shape(images, images, padding=padding, axes=[1, 2])
        spec = sfft.Fourier(images).fft(box, (1, 2))                                     # [C, Fy, Fx/2+1]
        ramp_y, ramp_x = mk_shifter(box)
        move = (ramp_y[None, :, None] ** (-np.asarray(y_hr, dtype=np.float64) * cos / hy)[:, None, None] *
                ramp_x[None, None, :] ** (-np.asarray(y_hr, dtype=np.float64) * sin / hx)[:, None, None])
        moved = sfft.Fourier.from_fft(spec[:, None] * move[None], box,
                                      (len(images), len(y_hr)) + images.shape[1:], [2, 3]).image    # [C, a, p, q]
        wy = np.sinc((np.asarray(y_lr)[None, :] + np.asarray(x_hr)[:, None] * sin) / hy)            # [b, p]
        wx = np.sinc((np.asarray(x_lr)[None, :] - np.asarray(x_hr)[:, None] * cos) / hx)            # [b, q]
        return np.einsum("capq,bp,bq->cab", moved, wy, wx)
    wy = np.sinc((y_lr[None, :] - y_hr[:, None]) / hy)
    wx = np.sinc((x_lr[:, None] - x_hr[None, :]) / hx)
    return np.array([wy @ img.T @ wx for img in images])


# ---------------------------------------------------------------- sub-pixel shifts (update.translation)
# 1-D interpolation kernels sampled at the integer offsets of their window: a handful of taps per axis, evaluated on
# the host (they are parameters of the shift, like a step size); the image work is the separable convolution
# `scl_resample_f32` on the device.
def _check_shift(dx):
    if np.abs(dx) > 1:
        raise ValueError("The fractional shift dx must be between -1 and 1")


def bilinear(dx):
    """Two taps weighting the neighbouring pixels by the distance to them (interpolation.py:139-165)."""
    _check_shift(dx)
    f = abs(dx)
    if dx >= 0:
        return np.array([1 - f, f]), np.arange(2)
    return np.array([f, 1 - f]), np.array([-1, 0])


def cubic_spline(dx, a=1, b=0):
    """Four taps of the BC-spline family (Mitchell & Netravali 1988 with B = b, C = a) around floor(dx)
    (interpolation.py:168-213)."""
    _check_shift(dx)
    window = np.arange(-1, 3) + np.floor(dx)
    x = np.abs(dx - window)
    # coefficients of 1, x, x^2, x^3 (times 6) on |x| <= 1 and 1 < |x| < 2
    inner = np.array([6 - 2 * b, 0, 6 * a + 12 * b - 18, 12 - 6 * a - 9 * b], dtype=float)
    outer = np.array([24 * a + 8 * b, -48 * a - 12 * b, 30 * a + 6 * b, -6 * a - b], dtype=float)
    powers = x[:, None] ** np.arange(4)[None, :]
    y = np.where(x <= 1, powers @ inner, np.where(x < 2, powers @ outer, 0.0)) / 6
    return y, window.astype(int)


def catmull_rom(dx):
    """BC-spline with a = 1/2, b = 0 (interpolation.py:216-221)."""
    return cubic_spline(dx, a=.5, b=0)


def mitchel_netravali(dx):
    """BC-spline with a = b = 1/3 (interpolation.py:224-230)."""
    return cubic_spline(dx, a=1 / 3, b=1 / 3)


def lanczos(dx, a=3):
    """2a taps of sinc(x) sinc(x / a) around floor(dx) (interpolation.py:233-252)."""
    _check_shift(dx)
    window = np.arange(-a + 1, a + 1) + np.floor(dx)
    x = dx - window
    return np.sinc(x) * np.sinc(x / a), window.astype(int)


def quintic_spline(dx, dtype=np.float64):
    """Seven taps of the piecewise quintic of interpolation.py:255-270 on the fixed window -3 .. 3."""
    window = np.arange(-3, 4)
    x = np.abs(dx - window)
    inner = 1 + x ** 3 / 12 * (-95 + 138 * x - 55 * x ** 2)
    middle = (x - 1) * (x - 2) / 24 * (-138 + 348 * x - 249 * x ** 2 + 55 * x ** 3)
    outer = (x - 2) * (x - 3) ** 2 / 24 * (-54 + 50 * x - 11 * x ** 2)
    return np.where(x <= 1, inner, np.where(x <= 2, middle, np.where(x <= 3, outer, 0.0))), window


def get_separable_kernel(dy, dx, kernel=lanczos, **kwargs):
    """Outer product of the two 1-D kernels and their windows (interpolation.py:273-299)."""
    kx, x_window = kernel(dx, **kwargs)
    ky, y_window = kernel(dy, **kwargs)
    return np.outer(ky, kx), y_window, x_window


def fft_resample(img, dy, dx, kernel=lanczos, **kwargs):
    """`img` shifted by the fraction of a pixel (dy, dx): convolution with the separable interpolation kernel, zero
    outside the image (interpolation.py:408-448).  The reference multiplies in Fourier space on a box padded by the
    kernel; a direct sum over the 2-7 taps per axis on the device is the same linear convolution."""
    from . import _native
    ky, ywin = kernel(dy, **kwargs)
    kx, xwin = kernel(dx, **kwargs)
    img = np.asarray(img)
    out = _native.resample(np.ascontiguousarray(img, dtype=np.float32)[None], [ky], [int(ywin[0])], [kx], [int(xwin[0])])[0]
    return out.astype(img.dtype if img.dtype.kind == "f" else np.float64, copy=False)
