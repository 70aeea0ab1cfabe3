"""PSF model images (host-side setup code; mirrors scarlet/psf.py:8-89 and
interpolation.apply_2D_trapezoid_rule, interpolation.py:497-552)."""
from functools import partial

import numpy as np

from .fft import Fourier


def moffat(y, x, y0, x0, amplitude, alpha, beta=1.5):
    X, Y = np.meshgrid(x, y)
    return amplitude * (1 + ((X - x0) ** 2 + (Y - y0) ** 2) / alpha ** 2) ** -beta


def gaussian(y, x, y0=0, x0=0, amplitude=None, sigma=1):
    if amplitude is None:
        amplitude = 1 / (np.pi ** 2 * sigma ** 2)
    X, Y = np.meshgrid(x, y)
    return amplitude * np.exp(-((X - x0) ** 2 + (Y - y0) ** 2) / (2 * sigma ** 2))


def double_gaussian(y, x, y0=0, x0=0, A1=None, sigma1=1, A2=None, sigma2=1):
    return gaussian(y, x, y0, x0, A1, sigma1) + gaussian(y, x, y0, x0, A2, sigma2)


def _trapezoid_image(y, x, f, n):
    dy, dx = y[1] - y[0], x[1] - x[0]
    _x = np.linspace(x[0] - dx / 2, x[-1] + dx / 2, len(x) * n + 1)
    _y = np.linspace(y[0] - dy / 2, y[-1] + dy / 2, len(y) * n + 1)
    z = f(_y, _x)
    dz = 0.4 * (z[:-1, :-1] + z[1:, :-1] + z[:-1, 1:] + z[1:, 1:])
    vol = dy * dx * dz / n / n
    rows, cols = len(_y) // n, len(_x) // n
    return np.array(np.split(np.array(np.split(vol, cols, axis=1)), rows, axis=1)).sum(axis=(2, 3))


def generate_psf_image(func, shape, subsamples=10, normalize=True, **kwargs):
    """Sub-pixel integrated PSF image wrapped in a `Fourier` (psf.py:55-89)."""
    ry, rx = np.array(shape) // 2
    y = np.linspace(-ry, ry, shape[0])
    x = np.linspace(-rx, rx, shape[1])
    result = _trapezoid_image(y, x, partial(func, **kwargs), subsamples)
    if normalize:
        result /= result.sum()
    return Fourier(result)
