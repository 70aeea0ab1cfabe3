"""Same-named stand-in for the reference's native module `scarlet.operators_pybind11`
(/root/reference/scarlet/operators_pybind11.cc:72-91): the three functions with the reference's argument lists, run on
the device through the C ABI (scl_pb_* / scl_apply_filter_*, include/scarlet_b200.h) on the CALLER's tables -- the
`weights`, `offsets`, `ref_idx` and `dist_idx` of operator.prox_strict_monotonic (operator.py:81-122), whatever order
they are in.  A user who swaps only the native module keeps `operator.py:24-37` working unchanged.

Binding behaviour kept from pybind11: `flat_img` / `X` are updated in place and must be writeable 1-D arrays of exactly
the overload's dtype (no-copy `Eigen::Ref` / `mutable_unchecked`: anything else is a TypeError); `weights` is cast to the
image dtype (operators_pybind11.cc:82-85); `step` is accepted and ignored; returns None.
"""
import ctypes

import numpy as np

from . import _native

__all__ = ["prox_monotonic", "prox_weighted_monotonic", "apply_filter"]

_c = ctypes.c_void_p


def _flat_inplace(a, name, dtypes):
    if not isinstance(a, np.ndarray) or a.ndim != 1 or a.dtype not in dtypes or not a.flags.writeable:
        raise TypeError("%s(): incompatible function arguments: a writeable 1-D %s array is needed" %
                        (name, " or ".join(np.dtype(d).name for d in dtypes)))
    return a


def _index(v, name):
    a = np.asarray(v)
    if a.dtype.kind not in "iu" or a.ndim != 1:
        if a.size == 0:
            return np.zeros(0, np.int32)
        raise TypeError("%s must be a 1-D sequence of integers" % name)
    return np.ascontiguousarray(a, dtype=np.int32)


def prox_weighted_monotonic(flat_img, step, weights, offsets, dist_idx, thresh):
    """operators_pybind11.cc:27-50: in `dist_idx` order, x[d] = min(x[d], (1 - thresh) * sum_i x[d + offsets[i]] *
    weights[i, d]) over the taps with a positive weight.  float32 and float64 overloads, chosen by flat_img.dtype."""
    _native.require_device()
    x = _flat_inplace(flat_img, "prox_weighted_monotonic", (np.float32, np.float64))
    w = np.ascontiguousarray(weights, dtype=x.dtype)
    off = _index(offsets, "offsets")
    didx = _index(dist_idx, "dist_idx")
    if w.ndim != 2 or w.shape != (len(off), x.size):
        raise TypeError("prox_weighted_monotonic(): weights must have shape (len(offsets), flat_img.size)")
    buf = x if x.flags.c_contiguous else np.ascontiguousarray(x)
    L = _native.lib()
    if x.dtype == np.float32:
        rc = L.scl_pb_prox_weighted_monotonic_f32(_native._p(buf), ctypes.c_longlong(x.size), _native._p(w), _native._p(off),
                                                  len(off), _native._p(didx), ctypes.c_longlong(len(didx)),
                                                  ctypes.c_float(thresh), None)
    else:
        rc = L.scl_pb_prox_weighted_monotonic_f64(_native._p(buf), ctypes.c_longlong(x.size), _native._p(w), _native._p(off),
                                                  len(off), _native._p(didx), ctypes.c_longlong(len(didx)),
                                                  ctypes.c_double(thresh), None)
    _native._check(rc)
    if buf is not x:
        x[...] = buf


def prox_monotonic(X, step, ref_idx, dist_idx, thresh):
    """operators_pybind11.cc:11-25: x[d] = min(x[d], x[ref_idx[d]] * (1 - thresh)) for d in dist_idx; float64 only."""
    _native.require_device()
    x = _flat_inplace(X, "prox_monotonic", (np.float64,))
    ref = _index(ref_idx, "ref_idx")
    didx = _index(dist_idx, "dist_idx")
    buf = x if x.flags.c_contiguous else np.ascontiguousarray(x)
    _native._check(_native.lib().scl_pb_prox_monotonic_f64(_native._p(buf), ctypes.c_longlong(x.size), _native._p(ref),
                                                           ctypes.c_longlong(len(ref)), _native._p(didx),
                                                           ctypes.c_longlong(len(didx)), ctypes.c_double(thresh), None))
    if buf is not x:
        x[...] = buf


def apply_filter(image, values, y_start, y_end, x_start, x_end, result):
    """operators_pybind11.cc:52-70 (no caller in the reference): result = sum_n values[n] * shifted block n of image;
    float32 and float64 overloads; `result` [rows, cols] is overwritten."""
    _native.require_device()
    if not isinstance(result, np.ndarray) or result.ndim != 2 or result.dtype not in (np.float32, np.float64) or \
            not result.flags.writeable:
        raise TypeError("apply_filter(): result must be a writeable 2-D float32 or float64 array")
    img = np.ascontiguousarray(image, dtype=result.dtype)
    if img.shape != result.shape:
        raise TypeError("apply_filter(): image and result shapes differ")
    vals = np.ascontiguousarray(values, dtype=result.dtype).reshape(-1)
    idx = [_index(v, n) for v, n in ((y_start, "y_start"), (y_end, "y_end"), (x_start, "x_start"), (x_end, "x_end"))]
    if any(len(i) != len(vals) for i in idx):
        raise TypeError("apply_filter(): values and the four index vectors must have one entry per filter tap")
    out = np.empty(result.shape, result.dtype)
    fn = _native.lib().scl_apply_filter_f32 if result.dtype == np.float32 else _native.lib().scl_apply_filter_f64
    _native._check(fn(_native._p(img), img.shape[0], img.shape[1], _native._p(vals), *[_native._p(i) for i in idx],
                      len(vals), _native._p(out), None))
    result[...] = out
