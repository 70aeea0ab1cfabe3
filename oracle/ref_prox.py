"""Oracle: array-level proximal operators and centre measurement
(TEST INFRASTRUCTURE ONLY).

Follows /root/reference/scarlet:
  sort_by_radius            operator.py:40-78
  neighbour_mask            operator.py:479-522 (diagonalizeArray edge masks)
  monotonic_weights         operator.py:540-621 (getRadialMonotonicWeights)
  prox_weighted_monotonic   operator.py:32-37,117-122 -> operators_pybind11.cc:27-50
  prox_nearest_monotonic    operator.py:24-29,105-115 -> operators_pybind11.cc:11-25
  uncentered_slices         operator.py:175-228 (uncentered_operator)
  prox_sdss/soft/kspace_symmetry   operator.py:231-288
  prox_uncentered_symmetry  operator.py:291-350
  mk_shifter                interpolation.py:302-340
  max_pixel, psf_weighted_centroid   measurement.py:3-94
  prox_plus                 proxmin.operators.prox_plus (third-party, absent; the
                            published definition X[X<0]=0 is pinned by the reference's
                            tests/test_update.py:23-44)
  normalized                update.py:35-68
"""
import ctypes
import os
import numpy as np
from . import ref_fft

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

COORDS = [(-1, -1), (-1, 0), (-1, 1), (0, -1), (0, 1), (1, -1), (1, 0), (1, 1)]


def _lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "_build", "liborc_monotonic.so")
        if not os.path.exists(path):
            import subprocess
            subprocess.check_call(["make", "-C", _HERE], stdout=subprocess.DEVNULL)
        _LIB = ctypes.CDLL(path)
    return _LIB


def sort_by_radius(shape, center=None, kind=None):
    """argsort of the Euclidean distance to `center` (operator.py:70-78).

    The reference uses numpy's default (unstable) argsort, whose tie order is
    numpy-version dependent; ties never influence the sweep.  The
    canonical order used everywhere in this repo is kind="stable" on the
    integer r^2 (ties by flat index)."""
    if center is None:
        cx = (shape[1] - 1) >> 1
        cy = (shape[0] - 1) >> 1
    else:
        cy, cx = int(center[0]), int(center[1])
    X, Y = np.meshgrid(np.arange(shape[1]) - cx, np.arange(shape[0]) - cy)
    if kind is None:
        return np.argsort(np.sqrt(X ** 2 + Y ** 2).flatten())
    return np.argsort((X ** 2 + Y ** 2).flatten(), kind="stable")


def radius2(shape, center):
    cy, cx = int(center[0]), int(center[1])
    X, Y = np.meshgrid(np.arange(shape[1]) - cx, np.arange(shape[0]) - cy)
    return (X ** 2 + Y ** 2).flatten()


def neighbour_mask(shape):
    """mask[i, p] = True where neighbour i of pixel p falls outside the image
    (the net effect of diagonalizeArray's slice + edge masks, operator.py:505-520)."""
    H, W = shape
    y, x = np.divmod(np.arange(H * W), W)
    mask = np.zeros((8, H * W), dtype=bool)
    for i, (oy, ox) in enumerate(COORDS):
        ny, nx = y + oy, x + ox
        mask[i] = (ny < 0) | (ny >= H) | (nx < 0) | (nx >= W)
    return mask


def monotonic_weights(shape, center=None, use_nearest=False):
    """(8, N) float64 weights of getRadialMonotonicWeights (operator.py:540-621)."""
    H, W = shape
    if center is None:
        center = ((H - 1) // 2, (W - 1) // 2)
    py, px = int(center[0]), int(center[1])
    X, Y = np.meshgrid(np.arange(W) - px, np.arange(H) - py)
    r2 = (X ** 2 + Y ** 2).flatten()
    inf = X == 0
    tX = X.copy()
    tX[inf] = 1
    angles = np.arctan2(-Y, -tX)
    sel = inf & (Y != 0)
    angles[sel] = 0.5 * np.pi * np.sign(angles[sel])
    angles = angles.flatten()
    mask = neighbour_mask(shape)
    yy, xx = np.divmod(np.arange(H * W), W)
    cosw = np.zeros((8, H * W), dtype=np.float64)
    for i, (oy, ox) in enumerate(COORDS):
        dy = np.full(H * W, float(oy))
        dx = np.full(H * W, float(ox))
        infn = dx == 0
        dx[infn] = 1
        rel = np.arctan2(dy, dx)
        s = infn & (dy != 0)
        rel[s] = 0.5 * np.pi * np.sign(rel[s])
        w = np.cos(angles - rel)
        # neighbour must be strictly closer to the peak (operator.py:569-571)
        ny = np.clip(yy + oy, 0, H - 1)
        nx = np.clip(xx + ox, 0, W - 1)
        nr2 = (ny - py) ** 2 + (nx - px) ** 2
        w[r2 - nr2 <= 0] = 0
        w[mask[i]] = 0
        cosw[i] = w
    if use_nearest:
        out = np.zeros_like(cosw)
        out[np.argmax(cosw, axis=0), np.arange(H * W)] = 1
        out[:, px + py * W] = 0
        return out
    norm = np.sum(cosw, axis=0)
    norm[norm == 0] = 1
    out = (cosw.T / norm[:, None]).T
    out[mask] = 0
    return out


def offsets(width):
    return np.array([width * y + x for y, x in COORDS], dtype=np.int32)


def prox_weighted_monotonic(X, weights, didx, offs, thresh=0):
    """In-place sweep; dtype of X selects the f32/f64 overload and the weights
    are cast to it first (operators_pybind11.cc:82-85)."""
    flat = X.reshape(-1)
    T = X.dtype
    w = np.ascontiguousarray(weights, dtype=T)
    didx = np.ascontiguousarray(didx, dtype=np.int32)
    offs = np.ascontiguousarray(offs, dtype=np.int32)
    lib = _lib()
    if T == np.float32:
        fn, ct = lib.orc_prox_weighted_monotonic_f32, ctypes.c_float
    elif T == np.float64:
        fn, ct = lib.orc_prox_weighted_monotonic_f64, ctypes.c_double
    else:
        raise TypeError("prox_weighted_monotonic: float32/float64 only")
    buf = np.ascontiguousarray(flat)
    fn(buf.ctypes.data_as(ctypes.c_void_p), w.ctypes.data_as(ctypes.c_void_p),
       ctypes.c_long(w.shape[1]), offs.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(len(offs)),
       didx.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(len(didx)), ct(thresh))
    flat[:] = buf
    return X


def prox_weighted_monotonic_py(X, weights, didx, offs, thresh=0):
    """Pure-Python twin of the C sweep (small cases; cross-checks the C build)."""
    flat = X.reshape(-1)
    T = X.dtype.type
    w = np.asarray(weights, dtype=X.dtype)
    for p in didx:
        ref = T(0)
        for i in range(len(offs)):
            if w[i, p] > 0:
                ref = T(ref + T(flat[offs[i] + p] * w[i, p]))
        r = T(ref * T(T(1) - T(thresh)))
        flat[p] = min(flat[p], r) if not r < flat[p] else r
    return X


def prox_nearest_monotonic(X, ref_idx, didx, thresh=0):
    flat = np.ascontiguousarray(X.reshape(-1), dtype=np.float64)
    ref_idx = np.ascontiguousarray(ref_idx, dtype=np.int32)
    didx = np.ascontiguousarray(didx, dtype=np.int32)
    _lib().orc_prox_monotonic_f64(flat.ctypes.data_as(ctypes.c_void_p),
                                  ref_idx.ctypes.data_as(ctypes.c_void_p),
                                  didx.ctypes.data_as(ctypes.c_void_p),
                                  ctypes.c_int(len(didx)), ctypes.c_double(thresh))
    X.reshape(-1)[:] = flat
    return X


def nearest_ref_idx(shape):
    """ref_idx of the use_nearest operator (operator.py:105-115): for each pixel
    the flat index of its single reference pixel; the peak refers to itself
    (getRadialMonotonicOp's -1 diagonal entry, operator.py:641-647)."""
    H, W = shape
    w = monotonic_weights(shape, None, use_nearest=True)
    offs = offsets(W)
    ref = np.arange(H * W)
    tap = np.argmax(w, axis=0)
    has = w.max(axis=0) > 0
    ref[has] = (np.arange(H * W) + offs[tap])[has]
    return ref


def build_monotonic(shape, center=None, thresh=0, kind="stable"):
    """The operator built by prox_strict_monotonic(use_nearest=False) (operator.py:116-122)."""
    didx = sort_by_radius(shape, center, kind=kind)
    w = monotonic_weights(shape, center)
    offs = offsets(shape[1])

    def prox(X, step=0):
        return prox_weighted_monotonic(X, w, didx[1:], offs, thresh)
    return prox


# --------------------------------------------------------------------------
# symmetry
# --------------------------------------------------------------------------
def uncentered_slices(shape, center):
    """Largest window of `shape` in which `center` is the //2 centre
    (operator.py:197-220).  Returns None when `center` is the frame centre
    (the reference then calls func on the whole array and RETURNS the result,
    operator.py:203-204)."""
    py, px = center
    cy, cx = shape[0] // 2, shape[1] // 2
    if py == cy and px == cx:
        return None
    dy = int(2 * (py - cy))
    dx = int(2 * (px - cx))
    if not shape[0] % 2:
        dy += 1
    if not shape[1] % 2:
        dx += 1
    xs = slice(None, dx) if dx < 0 else slice(dx, None)
    ys = slice(None, dy) if dy < 0 else slice(dy, None)
    return ys, xs


def uncentered_operator(X, func, center=None, fill=None, **kwargs):
    if center is None:
        center = np.unravel_index(np.argmax(X), X.shape)
    sl = uncentered_slices(X.shape, center)
    if sl is None:
        return func(X, **kwargs)
    if fill is not None:
        _X = np.ones(X.shape, X.dtype) * fill
        _X[sl] = func(X[sl], **kwargs)
        X[:] = _X
    else:
        X[sl] = func(X[sl], **kwargs)
    return X


def prox_sdss_symmetry(X, step=0):
    Xs = np.fliplr(np.flipud(X))
    X[:] = np.min([X, Xs], axis=0)
    return X


def prox_soft_symmetry(X, step=0, strength=1):
    Xs = np.fliplr(np.flipud(X))
    X[:] = 0.5 * strength * (X + Xs) + (1 - strength) * X
    return X


def mk_shifter(shape):
    """interpolation.py:326-334 (real=False)."""
    fy = np.fft.fftfreq(shape[0])
    fx = np.fft.rfftfreq(shape[1])
    return np.exp(-1j * 2 * np.pi * fy), np.exp(-1j * 2 * np.pi * fx)


def prox_kspace_symmetry(X, step=0, shift=None, padding=10):
    """operator.py:253-288.  NOT in place: returns a new array."""
    shape = ref_fft.fft_shape(X.shape, X.shape, padding=padding)
    dy, dx = shift
    X_fft = ref_fft.forward(X, shape, (0, 1))
    zero = X <= 0
    sy, sx = mk_shifter(shape)
    res = X_fft * sy[:, None] ** (-dy)
    res = res * sx[None, :] ** (-dx)
    res = res.real
    res = res * sy[:, None] ** dy
    res = res * sx[None, :] ** dx
    img = ref_fft.inverse(res, shape, X.shape, (0, 1))
    img[zero] = 0
    return np.real(img)


def prox_uncentered_symmetry(X, step=0, center=None, algorithm="kspace", fill=None,
                             shift=None, strength=.5):
    """operator.py:291-350, including the tuple-vs-0 comparison quirk
    (`np.all((dy, dx) == 0)` is False for a tuple)."""
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


# --------------------------------------------------------------------------
# centre measurement (measurement.py)
# --------------------------------------------------------------------------
def max_pixel(morph, center=None, window=None):
    if center is None:
        center = (morph.shape[0] // 2, morph.shape[1] // 2)
    cy, cx = int(center[0]), int(center[1])
    if window is None:
        window = slice(cy - 2, cy + 3), slice(cx - 2, cx + 3)
    _m = morph[window]
    yx0 = np.array([window[0].start, window[1].start])
    return tuple(np.unravel_index(np.argmax(_m), _m.shape) + yx0)


def psf_weighted_centroid(morph, psf, pixel_center):
    cy, cx = pixel_center
    rad = psf.shape[1] // 2
    ye = np.arange(morph.shape[0]) - cy
    xe = np.arange(morph.shape[1]) - cx
    ry = np.min((abs(ye[0]), abs(ye[-1]), rad))
    rx = np.min((abs(xe[0]), abs(xe[-1]), rad))
    ty = ye[abs(ye) <= ry]
    tx = xe[abs(xe) <= rx]
    mv = morph[ty + cy][:, tx + cx]
    pv = psf[ty + rad][:, tx + rad]
    w = mv * pv
    wsum = np.sum(w)
    indy, indx = np.indices(pv.shape)
    my = np.sum(indy * w) / wsum
    mx = np.sum(indx * w) / wsum
    offset = ((ty + cy)[0], (tx + cx)[0])
    whole = np.round((my, mx))
    dy, dx = whole - (my, mx)
    return tuple((whole + offset).astype(int)), (dy, dx)


# --------------------------------------------------------------------------
# positivity / normalisation (update.py:27-68)
# --------------------------------------------------------------------------
def prox_plus(X, step=0):
    X[X < 0] = 0
    return X


def normalized(sed, morph, type="morph_max"):
    t = type.lower()
    if t == "sed":
        norm = sed.sum()
        sed[:] = sed / norm
        morph[:] = morph * norm
    elif t == "morph":
        norm = morph.sum()
        sed[:] = sed * norm
        morph[:] = morph / norm
    elif t == "morph_max":
        norm = morph.max()
        sed[:] = sed * norm
        morph[:] = morph / norm
    else:
        raise ValueError("Unrecognized normalization '{0}'".format(type))


# --------------------------------------------------------------------------
# constraints the reference exports for custom update() schedules (update.py:71-103)
# --------------------------------------------------------------------------
def prox_hard(X, step, thresh=0):
    """proxmin.operators.prox_hard (third-party, absent; proxmin 0.6 operators.py: X[|X| < thresh * step] = 0),
    pinned by the reference's tests/test_update.py:74-86; caller update.sparse_l0 (update.py:71-75)."""
    X[np.abs(X) < thresh * step] = 0
    return X


def prox_soft(X, step, thresh=0):
    """proxmin.operators.prox_soft: sign(X) * max(|X| - thresh * step, 0); pinned by tests/test_update.py:88-97;
    caller update.sparse_l1 (update.py:78-82)."""
    X[:] = np.sign(X) * np.maximum(np.abs(X) - thresh * step, 0)
    return X


def threshold(morph):
    """measurement.threshold (measurement.py:96-113): cut at the last empty bin of the log10 histogram."""
    _morph = morph[morph > 0]
    _bins = 50
    if _morph.size < 500:
        _bins = max(int(_morph.size / 10), 1)
        if _bins == 1:
            return 0, _bins
    hist, bins = np.histogram(np.log10(_morph).reshape(-1), _bins)
    cutoff = np.where(hist == 0)[0]
    if len(cutoff) == 0:
        return 0, _bins
    return 10 ** bins[cutoff[-1]], _bins


def update_threshold(morph):
    """update.threshold (update.py:85-103) without the bounding box: pixels below the cut are zeroed."""
    thresh, _ = threshold(morph)
    morph[morph < thresh] = 0
    return morph
