"""Array-level proximal operators (the names of scarlet/operator.py), each
dispatching to a hand-written CUDA kernel through the C ABI.  Like the reference
they mutate `X` in place and return it; `step` is accepted and ignored
(operator.py:17-350).  There is no CPU path.
"""
from functools import partial

import numpy as np

from . import _native


def _as3(X):
    if X.ndim != 2:
        raise ValueError("expected a 2-D image")
    return X


def prox_max_unity(X, step):
    """operator.py:17-21 (trivial; host arithmetic on the caller's array)."""
    X[:] = X / X.max()
    return X


def prox_center_on(X, step, tiny=1e-10):
    """operator.py:149-158"""
    cy, cx = X.shape[0] // 2, X.shape[1] // 2
    X[cy, cx] = max(X[cy, cx], tiny)
    return X


def prox_sed_on(X, step, tiny=1e-10):
    """operator.py:161-172"""
    if np.all(X <= 0):
        X[:] = tiny
    return X


def sort_by_radius(shape, center=None):
    """Pixel indices sorted by distance from `center` (operator.py:40-78).  Canonical order:
    stable sort on the integer r^2 (ties by flat index); the reference's tie order is
    numpy-version dependent and never affects results."""
    if center is None:
        cx, cy = (shape[1] - 1) >> 1, (shape[0] - 1) >> 1
    else:
        cy, cx = int(center[0]), int(center[1])
    X, Y = np.meshgrid(np.arange(shape[1]) - cx, np.arange(shape[0]) - cy)
    return np.argsort((X ** 2 + Y ** 2).flatten(), kind="stable")


def _run_inplace(X, fn):
    """Run a batched device operator on one host image of any float dtype, in place."""
    buf = np.ascontiguousarray(X)[None].copy()
    fn(buf)
    X[...] = buf[0]
    return X


def _prox_weighted_monotonic(X, step, center, thresh=0):
    """operators_pybind11.prox_weighted_monotonic through the device wavefront (operator.py:32-37)."""
    X2 = X.reshape(X.shape[-2:]) if X.ndim == 2 else X
    if X.dtype not in (np.float32, np.float64):
        raise TypeError("prox_weighted_monotonic needs a float32 or float64 image")
    return _run_inplace(X2, lambda b: _native.prox_weighted_monotonic(b, [center], thresh))


def _prox_strict_monotonic(X, step, center, thresh=0):
    """operators_pybind11.prox_monotonic (nearest reference pixel, float64; operator.py:24-29)."""
    if X.dtype != np.float64:
        raise TypeError("prox_monotonic needs a float64 image")
    return _run_inplace(X, lambda b: _native.prox_monotonic_nearest(b, [center], thresh))


def prox_strict_monotonic(shape, use_nearest=False, thresh=0, center=None):
    """Build the monotonicity operator (operator.py:81-122).  The returned callable has the
    reference's signature `prox(X, step)`; ordering and weights are generated on the device."""
    height, width = shape
    if center is None:
        center = ((height - 1) >> 1, (width - 1) >> 1)
    center = (int(center[0]), int(center[1]))
    if use_nearest:
        if thresh != 0:
            raise ValueError("Thresholding does not work with nearest neighbor monotonicity")
        return partial(_prox_strict_monotonic, center=center, thresh=thresh)
    return partial(_prox_weighted_monotonic, center=center, thresh=thresh)


def uncentered_operator(X, func, center=None, fill=None, **kwargs):
    """Apply `func` on the largest window centred on `center` (operator.py:175-228)."""
    if center is None:
        py, px = np.unravel_index(np.argmax(X), X.shape)
    else:
        py, px = center
    cy, cx = np.array(X.shape) // 2
    if py == cy and px == cx:
        return func(X, **kwargs)
    dy, dx = int(2 * (py - cy)), int(2 * (px - cx))
    if not X.shape[0] % 2:
        dy += 1
    if not X.shape[1] % 2:
        dx += 1
    xs = slice(None, dx) if dx < 0 else slice(dx, None)
    ys = slice(None, dy) if dy < 0 else slice(dy, None)
    if fill is not None:
        _X = np.ones(X.shape, X.dtype) * fill
        _X[ys, xs] = func(X[ys, xs], **kwargs)
        X[:] = _X
    else:
        X[ys, xs] = func(X[ys, xs], **kwargs)
    return X


def _device_symmetry(X, mode, shift=None, strength=1.0):
    c = (-1, -1)
    buf = np.ascontiguousarray(X, dtype=np.float32)[None].copy()
    _native.prox_symmetry(buf, [c], None if shift is None else [shift], mode, strength)
    return buf[0].astype(X.dtype, copy=False)


def prox_sdss_symmetry(X, step):
    """min(X, flip(X)) (operator.py:231-239)"""
    X[:] = _device_symmetry(X, "sdss")
    return X


def prox_soft_symmetry(X, step, strength=1):
    """operator.py:242-251"""
    X[:] = _device_symmetry(X, "soft", None, strength)
    return X


def prox_kspace_symmetry(X, step, shift=None, padding=10):
    """Symmetry about a fractional centre (operator.py:253-288), evaluated on the device as the equivalent
    pair of Dirichlet-kernel operators.  Not in place: returns a new array, like the reference.
    """
    if padding != 10:
        raise NotImplementedError("only the reference's default padding=10 is supported")
    return _kspace_full(X, shift)


def _kspace_full(X, shift):
    buf = np.ascontiguousarray(X, dtype=np.float32)[None].copy()
    _native.prox_symmetry(buf, [(-1, -1)], [shift], "kspace", 1.0)      # (-1,-1): whole-array window
    return buf[0].astype(X.dtype, copy=False)


def prox_uncentered_symmetry(X, step, center=None, algorithm="kspace", fill=None, shift=None, strength=.5):
    """Symmetry with an off-centre peak (operator.py:291-350)."""
    if algorithm == "kspace" and (shift is None or np.all(shift == 0)):
        algorithm = "soft"
        strength = 1
    if algorithm == "kspace":
        return uncentered_operator(X, prox_kspace_symmetry, center, shift=shift, step=step, fill=fill)
    if algorithm == "sdss":
        return uncentered_operator(X, prox_sdss_symmetry, center, step=step, fill=fill)
    if algorithm == "soft":
        return uncentered_operator(X, prox_soft_symmetry, center, step=step, strength=strength, fill=fill)
    raise ValueError("algorithm must be one of 'soft', 'sdss', 'kspace', recieved '{0}''".format(algorithm))
