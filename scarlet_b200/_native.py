"""ctypes binding of libscarlet_b200.so (include/scarlet_b200.h) and the small
host-side tables the kernels consume.  There is no CPU fallback: every compute
entry point raises when the library or a CUDA device is missing.
"""
import ctypes
import os
import threading

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SCL_LIB") or os.path.join(_HERE, "lib", "libscarlet_b200.so")     # SCL_LIB: experiment builds

FFT_EDGE = 128
SPEC_COLS = 65

F_SYMMETRIC, F_MONOTONIC, F_FIX_SED, F_FIX_MORPH, F_KSPACE, F_CONSTRAINED, F_SHARED_CENTER, F_SCHEDULE = 1, 2, 4, 8, 16, 32, 64, 128
SCHED_OPS = 16      # entries of a declared schedule (include/scarlet_b200.h: SCL_SCHED_OPS)

_lib = None
_lock = threading.Lock()

c_int, c_float, c_double, c_void_p = ctypes.c_int, ctypes.c_float, ctypes.c_double, ctypes.c_void_p


class BatchDesc(ctypes.Structure):
    _fields_ = [
        ("n_blends", ctypes.c_int32), ("n_comp", ctypes.c_int32),
        ("C", ctypes.c_int32), ("Ny", ctypes.c_int32), ("Nx", ctypes.c_int32),
        ("comp_off", c_void_p), ("sed", c_void_p), ("morph", c_void_p), ("cflags", c_void_p),
        ("center", c_void_p), ("shift", c_void_p), ("conv", c_void_p),
        ("images", c_void_p), ("weights", c_void_p),
        ("n_ker", ctypes.c_int32), ("kernel_spec", c_void_p), ("ker_idx", c_void_p),
        ("n_cw", ctypes.c_int32), ("ph", ctypes.c_int32), ("pw", ctypes.c_int32),
        ("cw", c_void_p), ("cw_idx", c_void_p),
        ("it0", c_void_p), ("prev_loss", c_void_p), ("src_first", c_void_p),
        ("n_planes", ctypes.c_int32), ("plane_band", c_void_p), ("n_obs", ctypes.c_int32),
        ("n_lr_planes", ctypes.c_int32), ("lr_n", ctypes.c_int32), ("lr_images", c_void_p), ("lr_weights", c_void_p),
        ("n_lr_op", ctypes.c_int32), ("lr_op", c_void_p), ("lr_op_idx", c_void_p),
        ("ker_half", ctypes.c_int32), ("scene_mode", ctypes.c_int32),
        ("psf_obs", c_void_p), ("psf_model", c_void_p),
        ("psf_sets", ctypes.c_int32), ("psf_model_sets", ctypes.c_int32), ("psf_ky", ctypes.c_int32),
        ("psf_kx", ctypes.c_int32), ("psf_kym", ctypes.c_int32), ("psf_kxm", ctypes.c_int32),
        ("sched_ops", c_void_p), ("sched_args", c_void_p),
        ("prior_par", c_void_p), ("prior_b_sed", c_void_p), ("prior_idx", c_void_p), ("prior_b_morph", c_void_p),
        ("n_prior_morph", ctypes.c_int32),
    ]


def lib():
    """Load the CUDA library; fail loudly if it has not been built."""
    global _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise RuntimeError("libscarlet_b200.so is not built (run `python -m scarlet_b200.build`); "
                                   "scarlet_b200 has no CPU fallback")
            L = ctypes.CDLL(LIB_PATH)
            L.scl_last_error.restype = ctypes.c_char_p
            L.scl_batch_destroy.restype = None
            _lib = L
    return _lib


def _check(rc):
    if rc == 0:
        return
    msg = lib().scl_last_error().decode()
    if rc == -1:
        raise ValueError(msg)
    if rc == -3:
        raise NotImplementedError(msg)
    raise RuntimeError(msg)


def require_device():
    if lib().scl_device_count() < 1:
        raise RuntimeError("scarlet_b200 needs a CUDA device (no CPU fallback)")


def device_get():
    """CUDA device of the calling host thread."""
    d = c_int(0)
    _check(lib().scl_device_get(ctypes.byref(d)))
    return d.value


def device_set(device):
    """Select the CUDA device of the calling host thread (worker threads start on device 0)."""
    _check(lib().scl_device_set(int(device)))


def stream_create():
    """A CUDA stream handle (int) for the `stream=` arguments of this module."""
    s = c_void_p()
    _check(lib().scl_stream_create(ctypes.byref(s)))
    return s.value


def stream_destroy(stream):
    if stream:
        _check(lib().scl_stream_destroy(c_void_p(stream)))


def _p(a):
    return None if a is None else a.ctypes.data_as(c_void_p)


# --------------------------------------------------------------------------- host tables
def next_fast_len(target):
    """Smallest 5-smooth integer >= target (what scipy.fftpack.next_fast_len returns, fft.py:99)."""
    target = int(target)
    if target <= 6:
        return target
    best = None
    p5 = 1
    while p5 < 2 * target:
        p35 = p5
        while p35 < 2 * target:
            n = p35
            while n < target:
                n *= 2
            best = n if best is None or n < best else best
            p35 *= 3
        p5 *= 5
    return best


def conv_fft_shape(n_img, n_ker, last):
    """The reference's FFT length for one axis of fft.convolve (fft.py:68-106, padding 3)."""
    f = next_fast_len(n_img + n_ker + 3)
    if last:
        while f % 2:
            f = next_fast_len(f + 1)
    return f


_COORDS = [(-1, -1), (-1, 0), (-1, 1), (0, -1), (0, 1), (1, -1), (1, 0), (1, 1)]
_tables = {}
_registered = set()


def monotonic_tables(Ny, Nx):
    """Build (once per shape) and register the monotonicity tables with the CUDA library."""
    key = (int(Ny), int(Nx))
    dkey = (device_get(),) + key            # the library keeps its tables per CUDA device
    if dkey not in _registered:
        wtab, ctab = monotonic_tables_host(Ny, Nx)
        _check(lib().scl_set_monotonic_tables(c_int(Ny), c_int(Nx), _p(wtab), _p(ctab)))
        _registered.add(dkey)
    return _tables[key]


def monotonic_tables_host(Ny, Nx):
    """Offset-indexed weight tables of the radial monotonicity operator.

    Same arithmetic as getRadialMonotonicWeights (operator.py:540-621) evaluated for
    every offset (Y, X) = pixel - peak instead of for one peak position:
      ctab[Y, X, i]  cos(theta_p - phi_i) if neighbour i is strictly closer to the peak, else 0
      wt9[bt, Y, X, i]  ctab normalised over the taps that stay inside the image for a pixel of border type bt
                        (0 = interior), rounded to float32 -- what the f32 overload multiplies with
    ctab is kept for the float64 / nearest variants, which renormalise on the device.
    """
    key = (int(Ny), int(Nx))
    if key in _tables:
        return _tables[key]
    Y, X = np.meshgrid(np.arange(-(Ny - 1), Ny), np.arange(-(Nx - 1), Nx), indexing="ij")
    inf = X == 0
    tX = X.copy()
    tX[inf] = 1
    ang = np.arctan2(-Y, -tX)
    sel = inf & (Y != 0)
    ang[sel] = 0.5 * np.pi * np.sign(ang[sel])
    ctab = np.zeros(Y.shape + (8,), dtype=np.float64)
    for i, (oy, ox) in enumerate(_COORDS):
        dy = np.full(Y.shape, float(oy))
        dx = np.full(Y.shape, float(ox))
        infn = dx == 0
        dx[infn] = 1
        rel = np.arctan2(dy, dx)
        s = infn & (dy != 0)
        rel[s] = 0.5 * np.pi * np.sign(rel[s])
        w = np.cos(ang - rel)
        closer = (Y + oy) ** 2 + (X + ox) ** 2 < Y ** 2 + X ** 2
        w[~closer] = 0
        ctab[..., i] = w
    # border types: bt = by*3 + bx, by/bx = 0 interior, 1 first row/column, 2 last row/column; taps that leave the
    # image are dropped before the normalisation (operator.py:505-520,614-617)
    wt9 = np.zeros((9,) + Y.shape + (8,), dtype=np.float32)
    for by in range(3):
        for bx in range(3):
            keep = np.array([not ((by == 1 and oy == -1) or (by == 2 and oy == 1) or
                                  (bx == 1 and ox == -1) or (bx == 2 and ox == 1)) for oy, ox in _COORDS])
            c = ctab * keep
            norm = np.zeros(Y.shape)
            for i in range(8):                      # sequential tap-order sum, like np.sum(axis=0) on (8, N)
                norm = norm + c[..., i]
            norm[norm == 0] = 1
            wt9[by * 3 + bx] = (c / norm[..., None]).astype(np.float32)
    ctab = np.ascontiguousarray(ctab)
    wt9 = np.ascontiguousarray(wt9)
    _tables[key] = (wt9, ctab)
    return _tables[key]


def kernel_half(ky, kx, Ny, Nx):
    """Reach of a ky x kx difference kernel on either side of a pixel, as placed by kernel_spectrum."""
    Fy, Fx = conv_fft_shape(Ny, ky, False), conv_fft_shape(Nx, kx, True)
    offy = (Fy - ky + 1) // 2 - Fy // 2
    offx = (Fx - kx + 1) // 2 - Fx // 2
    return int(max(-offy, ky - 1 + offy, -offx, kx - 1 + offx))


def kernel_spectrum(kernels, Ny, Nx):
    """rfft2 of real-space difference kernels [n, ky, kx] on the 128x128 periodic grid.

    Tap j of an axis lands on grid index j + off with off = (F-k+1)//2 - F//2, F the
    reference's FFT length for that axis (fft.py:60-65,193-211) -- for odd kernels this is
    simply the centre tap at index 0.  Returns complex64 [n, 128, 65].
    """
    kernels = np.asarray(kernels)
    n, ky, kx = kernels.shape
    Fy, Fx = conv_fft_shape(Ny, ky, False), conv_fft_shape(Nx, kx, True)
    offy = (Fy - ky + 1) // 2 - Fy // 2
    offx = (Fx - kx + 1) // 2 - Fx // 2
    # frames that do not fit one tile together with the kernel's reach run as overlap-save tiles (scene mode)
    if 2 * kernel_half(ky, kx, Ny, Nx) >= FFT_EDGE - 8:
        raise NotImplementedError("difference kernel too wide for the 128-point convolution tile")
    grid = np.zeros((n, FFT_EDGE, FFT_EDGE), dtype=np.float64)
    iy = (np.arange(ky) + offy) % FFT_EDGE
    ix = (np.arange(kx) + offx) % FFT_EDGE
    grid[:, iy[:, None], ix[None, :]] = kernels
    return np.ascontiguousarray(np.fft.rfft2(grid).astype(np.complex64))


# --------------------------------------------------------------------------- operators
def _centers(centers, n):
    c = np.ascontiguousarray(np.asarray(centers, dtype=np.int32).reshape(n, 2))
    return c


def _inplace_stack(x, name, dtypes):
    """The in-place operators write through x's buffer: a sliced or transposed view would be read and written with
    the wrong strides."""
    if not isinstance(x, np.ndarray) or x.ndim != 3:
        raise TypeError("%s: a [n, Ny, Nx] numpy array expected" % name)
    if x.dtype not in dtypes:
        raise TypeError("%s: %s arrays only" % (name, " or ".join(np.dtype(d).name for d in dtypes)))
    if not x.flags.c_contiguous:
        raise TypeError("%s: C-contiguous arrays only (the operator works in place)" % name)


def prox_weighted_monotonic(x, centers, thresh=0.0):
    """In place on x [n, Ny, Nx] (float32 or float64, C-contiguous)."""
    require_device()
    _inplace_stack(x, "prox_weighted_monotonic", (np.float32, np.float64))
    n, Ny, Nx = x.shape
    monotonic_tables(Ny, Nx)
    c = _centers(centers, n)
    if x.dtype == np.float32:
        _check(lib().scl_prox_weighted_monotonic_f32(_p(x), n, Ny, Nx, _p(c), c_float(thresh), None))
    elif x.dtype == np.float64:
        _check(lib().scl_prox_weighted_monotonic_f64(_p(x), n, Ny, Nx, _p(c), c_double(thresh), None))
    else:
        raise TypeError("prox_weighted_monotonic: float32 or float64 arrays only")
    return x


def prox_monotonic_nearest(x, centers, thresh=0.0):
    require_device()
    _inplace_stack(x, "prox_monotonic", (np.float64,))
    n, Ny, Nx = x.shape
    monotonic_tables(Ny, Nx)
    _check(lib().scl_prox_monotonic_f64(_p(x), n, Ny, Nx, _p(_centers(centers, n)), c_double(thresh), None))
    return x


def prox_symmetry(x, centers, shifts=None, mode="kspace", strength=1.0):
    require_device()
    _inplace_stack(x, "prox_symmetry", (np.float32,))
    n, Ny, Nx = x.shape
    m = {"kspace": 0, "soft": 1, "sdss": 2}[mode]
    sh = None if shifts is None else np.ascontiguousarray(np.asarray(shifts, dtype=np.float64).reshape(n, 2))
    _check(lib().scl_prox_symmetry_f32(_p(x), n, Ny, Nx, _p(_centers(centers, n)), _p(sh), m, c_float(strength), None))
    return x


def prox_symmetry_noalloc(x_ptr, n, Ny, Nx, centers_ptr, shifts_ptr=None, mode="kspace", strength=1.0, n_bad_ptr=None,
                          stream=None):
    """scl_prox_symmetry_noalloc_f32: prox_symmetry on caller-owned DEVICE arrays given as integer addresses (e.g.
    torch `tensor.data_ptr()`): x float32 [n, Ny, Nx] in place, centers int32 [n, 2], shifts float64 [n, 2], n_bad
    int32 [1] (zeroed by the caller).  Stream-ordered, no copies, no device allocation: scratch is a static device
    array of 16 MiB (NotImplementedError beyond it).  No counterpart in the reference (include/scarlet_b200.h)."""
    require_device()
    m = {"kspace": 0, "soft": 1, "sdss": 2}[mode]
    _check(lib().scl_prox_symmetry_noalloc_f32(c_void_p(x_ptr), int(n), int(Ny), int(Nx), c_void_p(centers_ptr),
                                               c_void_p(shifts_ptr), m, c_float(strength), c_void_p(n_bad_ptr),
                                               c_void_p(stream)))


def measure(x, centers, psf=None, do_max=True, do_centroid=True):
    """max_pixel and/or psf_weighted_centroid for x [n, Ny, Nx] float32.
    Returns (centers int32 [n,2], shifts float64 [n,2])."""
    require_device()
    n, Ny, Nx = x.shape
    x = np.ascontiguousarray(x, dtype=np.float32)
    c = _centers(centers, n).copy()
    s = np.zeros((n, 2), dtype=np.float64)
    ph = pw = 0
    if psf is not None:
        psf = np.ascontiguousarray(psf, dtype=np.float32)
        ph, pw = psf.shape
    _check(lib().scl_measure_f32(_p(x), n, Ny, Nx, _p(c), _p(s), _p(psf), ph, pw, int(do_max), int(do_centroid), None))
    return c, s


def convolve(model, spec, ker_idx, adjoint=False, ker_half=0):
    """model float32 [n, C, Ny, Nx]; spec complex64 [n_ker,128,65]; ker_idx int32 [n*C]; ker_half: reach of the
    kernels (kernel_half), needed for frames beyond one convolution tile."""
    require_device()
    model = np.ascontiguousarray(model, dtype=np.float32)
    n, C, Ny, Nx = model.shape
    out = np.empty_like(model)
    ker_idx = np.ascontiguousarray(ker_idx, dtype=np.int32).reshape(-1)
    n_ker = 0 if spec is None else len(spec)
    if spec is not None:
        spec = np.ascontiguousarray(spec, dtype=np.complex64)
    _check(lib().scl_convolve_f32(_p(model), _p(out), n, C, Ny, Nx, _p(spec), n_ker, _p(ker_idx), int(adjoint),
                                  int(ker_half), None))
    return out


def lowres_render(model, op):
    """LowResObservation._render as its matrix: model float32 [n, C, N]; op float32 [C, n_lr, N] -> [n, C, n_lr]."""
    require_device()
    model = np.ascontiguousarray(model, dtype=np.float32)
    op = np.ascontiguousarray(op, dtype=np.float32)
    n, C, N = model.shape
    assert op.shape[0] == C and op.shape[2] == N
    out = np.empty((n, C, op.shape[1]), np.float32)
    _check(lib().scl_lowres_render_f32(_p(model), _p(out), n, C, N, op.shape[1], _p(op), None))
    return out


def match_psfs(psf_obs, psf_model, padding=3, frame_shape=None, want_diff=True):
    """fft.match_psfs for n PSF sets at once on the device (fft.py:282-301, observation.py:155-196).

    psf_obs float32 [n, C, ky, kx]; psf_model float32 [1 or n, kym, kxm].  Returns (diff [n, C, ky, kx] or None,
    spec complex64 [n, C, 128, 65] or None); the spectra are built when `frame_shape` = (Ny, Nx) is given.
    """
    require_device()
    psf_obs = np.ascontiguousarray(psf_obs, dtype=np.float32)
    psf_model = np.ascontiguousarray(psf_model, dtype=np.float32)
    if psf_obs.ndim != 4 or psf_model.ndim != 3:
        raise ValueError("match_psfs: psf_obs [n, C, ky, kx] and psf_model [1 or n, kym, kxm] expected")
    n, C, ky, kx = psf_obs.shape
    diff = np.empty((n, C, ky, kx), np.float32) if want_diff else None
    spec = np.empty((n, C, FFT_EDGE, FFT_EDGE // 2 + 1), np.complex64) if frame_shape is not None else None
    Ny, Nx = frame_shape if frame_shape is not None else (0, 0)
    _check(lib().scl_match_psfs_f32(_p(psf_obs), n, C, ky, kx, _p(psf_model), psf_model.shape[0], psf_model.shape[1],
                                    psf_model.shape[2], int(padding), _p(diff), _p(spec), int(Ny), int(Nx), None))
    return diff, spec


def trim(x, min_value=0.0):
    """bbox.trim / bbox.flux_at_edge / pixel sums of n images at once (bbox.py:170-210).

    x float32 [n, Ny, Nx] -> (bounds int32 [n, 4] = bottom, top, left, right or -1, edge bool [n], sum float64 [n]).
    """
    require_device()
    x = np.ascontiguousarray(x, dtype=np.float32)
    n, Ny, Nx = x.shape
    bounds = np.empty((n, 4), np.int32)
    edge = np.empty(n, np.int32)
    tot = np.empty(n, np.float64)
    _check(lib().scl_trim_f32(_p(x), n, Ny, Nx, c_float(min_value), _p(bounds), _p(edge), _p(tot), None))
    return bounds, edge.astype(bool), tot


def prox_sparse(x, thresh, mode):
    """In place on a float32 array: mode 0 hard threshold (prox_hard), 1 soft threshold (prox_soft)."""
    require_device()
    if x.dtype != np.float32 or not x.flags.c_contiguous:
        raise TypeError("prox_sparse: C-contiguous float32 arrays only")
    _check(lib().scl_prox_sparse_f32(_p(x), ctypes.c_longlong(x.size), c_float(thresh), int(mode), None))
    return x


def threshold(x, apply=False):
    """measurement.threshold of n images x float32 [n, Ny, Nx]: (thresh float64 [n], bins int32 [n]); with `apply`
    also update.threshold in place: -> (thresh, bins, bounds int32 [n, 4])."""
    require_device()
    if x.dtype != np.float32 or not x.flags.c_contiguous or x.ndim != 3:
        raise TypeError("threshold: C-contiguous float32 [n, Ny, Nx] arrays only")
    n, Ny, Nx = x.shape
    th = np.empty(n, np.float64)
    bins = np.empty(n, np.int32)
    bounds = np.empty((n, 4), np.int32) if apply else None
    _check(lib().scl_threshold_f32(_p(x), n, Ny, Nx, int(bool(apply)), _p(th), _p(bins), _p(bounds), None))
    return (th, bins, bounds) if apply else (th, bins)


def resample(x, ky, y0, kx, x0):
    """Separable interpolation-kernel convolution of n images (fft_resample): x float32 [n, Ny, Nx]; ky [n, Ly],
    kx [n, Lx] taps, y0 / x0 [n] offsets of the first tap.  Returns float32 [n, Ny, Nx]."""
    require_device()
    x = np.ascontiguousarray(x, dtype=np.float32)
    n, Ny, Nx = x.shape
    ky = np.ascontiguousarray(ky, dtype=np.float64).reshape(n, -1)
    kx = np.ascontiguousarray(kx, dtype=np.float64).reshape(n, -1)
    y0 = np.ascontiguousarray(y0, dtype=np.int32).reshape(n)
    x0 = np.ascontiguousarray(x0, dtype=np.int32).reshape(n)
    out = np.empty_like(x)
    _check(lib().scl_resample_f32(_p(x), _p(out), n, Ny, Nx, _p(ky), _p(y0), ky.shape[1], _p(kx), _p(x0), kx.shape[1],
                                  None))
    return out


class Fit64Desc(ctypes.Structure):
    _fields_ = [
        ("n_comp", ctypes.c_int32), ("C", ctypes.c_int32), ("Ny", ctypes.c_int32), ("Nx", ctypes.c_int32),
        ("sed", c_void_p), ("morph", c_void_p), ("cflags", c_void_p), ("center", c_void_p), ("shift", c_void_p),
        ("conv", c_void_p), ("src_first", c_void_p),
        ("n_planes", ctypes.c_int32), ("plane_band", c_void_p), ("n_obs", ctypes.c_int32),
        ("images", c_void_p), ("weights", c_void_p), ("ker_on", c_void_p), ("kernels", c_void_p),
        ("ky", ctypes.c_int32), ("kx", ctypes.c_int32), ("offy", ctypes.c_int32), ("offx", ctypes.c_int32),
        ("cw", c_void_p), ("ph", ctypes.c_int32), ("pw", ctypes.c_int32),
        ("it0", ctypes.c_int32), ("prev_loss", c_double),
    ]


def kernel_offsets(ky, kx, Ny, Nx):
    """Displacement of tap 0 of a ky x kx difference kernel on an Ny x Nx frame, as fft.convolve places it
    (fft.py:60-65,87-106): tap (j, i) acts at (j + offy, i + offx)."""
    Fy, Fx = conv_fft_shape(Ny, ky, False), conv_fft_shape(Nx, kx, True)
    return (Fy - ky + 1) // 2 - Fy // 2, (Fx - kx + 1) // 2 - Fx // 2


def _desc64(C, Ny, Nx, sed, morph, cflags, center, shift, conv, images, weights, plane_band, n_obs, kernels, ker_on,
            cw, it0, prev_loss, src_first):
    ca = np.ascontiguousarray
    N = Ny * Nx
    sed = ca(sed, dtype=np.float64)
    K = len(sed)
    plane_band = ca(plane_band, dtype=np.int32)
    P = len(plane_band)
    a = dict(sed=sed.reshape(K, C), morph=ca(morph, dtype=np.float64).reshape(K, N), cflags=ca(cflags, dtype=np.uint32).reshape(K),
             center=ca(center, dtype=np.int32).reshape(K, 2), shift=ca(shift, dtype=np.float64).reshape(K, 2),
             conv=ca(conv, dtype=np.int32).reshape(K), src_first=None if src_first is None else ca(src_first, dtype=np.int32),
             plane_band=plane_band, images=None if images is None else ca(images, dtype=np.float64).reshape(P, N),
             weights=None if weights is None else ca(weights, dtype=np.float64).reshape(P, N),
             ker_on=ca(ker_on, dtype=np.int32).reshape(P),
             kernels=None if kernels is None else ca(kernels, dtype=np.float64),
             cw=None if cw is None else ca(cw, dtype=np.float64))
    d = Fit64Desc()
    d.n_comp, d.C, d.Ny, d.Nx, d.n_planes, d.n_obs = K, C, Ny, Nx, P, int(n_obs)
    for name in ("sed", "morph", "cflags", "center", "shift", "conv", "src_first", "plane_band", "images", "weights",
                 "ker_on", "kernels", "cw"):
        setattr(d, name, _p(a[name]))
    if a["kernels"] is not None:
        assert a["kernels"].shape[0] == P
        d.ky, d.kx = a["kernels"].shape[1:]
        d.offy, d.offx = kernel_offsets(d.ky, d.kx, Ny, Nx)
    if a["cw"] is not None:
        d.ph, d.pw = a["cw"].shape
    d.it0, d.prev_loss = int(it0), float(prev_loss)
    return d, a, K


def constrain_f64(C, Ny, Nx, sed, morph, cflags, center, shift, cw, src_first=None):
    """scl_constrain_f64: the construction-time update() (it = 0) of float64 components; returns dict(sed, morph,
    center, shift)."""
    require_device()
    monotonic_tables(Ny, Nx)
    K = len(sed)
    d, a, K = _desc64(C, Ny, Nx, sed, morph, cflags, center, shift, np.zeros(K, np.int32), None, None, np.arange(C), 1,
                      None, np.zeros(C, np.int32), cw, 0, 0.0, src_first)
    out = dict(sed=np.empty((K, C)), morph=np.empty((K, Ny, Nx)), center=np.empty((K, 2), np.int32), shift=np.empty((K, 2)))
    _check(lib().scl_constrain_f64(ctypes.byref(d), _p(out["sed"]), _p(out["morph"]), _p(out["center"]), _p(out["shift"]),
                                   None))
    return out


def fit_f64(C, Ny, Nx, sed, morph, cflags, center, shift, conv, images, weights, plane_band, n_obs, kernels, ker_on,
            cw, it0=0, prev_loss=0.0, src_first=None, max_iter=200, e_rel=1e-2, approximate_L=False):
    """scl_fit_f64: Blend.fit of ONE blend in float64 (include/scarlet_b200.h).  kernels [P, ky, kx] or None."""
    require_device()
    monotonic_tables(Ny, Nx)
    d, a, K = _desc64(C, Ny, Nx, sed, morph, cflags, center, shift, conv, images, weights, plane_band, n_obs, kernels,
                      ker_on, cw, it0, prev_loss, src_first)
    out = dict(sed=np.empty((K, C)), morph=np.empty((K, Ny, Nx)), center=np.empty((K, 2), np.int32),
               shift=np.empty((K, 2)), conv=np.empty(K, np.int32), logL=np.zeros(max(int(max_iter), 1)),
               lipschitz=np.zeros(2))
    n_iter, done = ctypes.c_int32(0), ctypes.c_int32(0)
    _check(lib().scl_fit_f64(ctypes.byref(d), int(max_iter), c_double(e_rel), int(bool(approximate_L)), _p(out["sed"]),
                             _p(out["morph"]), _p(out["center"]), _p(out["shift"]), _p(out["conv"]), ctypes.byref(n_iter),
                             ctypes.byref(done), _p(out["logL"]), _p(out["lipschitz"]), None))
    out["n_iter"], out["converged"] = n_iter.value, bool(done.value)
    if n_iter.value == 0:
        out["sed"], out["morph"] = a["sed"].copy(), a["morph"].reshape(K, Ny, Nx).copy()
        out["center"], out["shift"], out["conv"] = a["center"].copy(), a["shift"].copy(), a["conv"].copy()
    out["logL"] = out["logL"][:n_iter.value]
    return out


def convolve_f64(model, kernels, adjoint=False):
    """Observation.render in float64: model [n, C, Ny, Nx], real-space difference kernels [C, ky, kx]."""
    require_device()
    model = np.ascontiguousarray(model, dtype=np.float64)
    kernels = np.ascontiguousarray(kernels, dtype=np.float64)
    n, C, Ny, Nx = model.shape
    ky, kx = kernels.shape[1:]
    offy, offx = kernel_offsets(ky, kx, Ny, Nx)
    out = np.empty_like(model)
    _check(lib().scl_convolve_f64(_p(model), _p(out), n, C, Ny, Nx, _p(kernels), ky, kx, offy, offx, int(adjoint), None))
    return out


# --------------------------------------------------------------------------- batch fit
class Batch:
    """Device-resident state of B independent blends sharing one frame shape
    (thin owner of a `scl_batch*`)."""

    def __init__(self, C, Ny, Nx, comp_off, sed, morph, cflags, center, shift, conv, images, weights,
                 kernel_spec, ker_idx, cw, cw_idx, it0=None, prev_loss=None, src_first=None, stream=None,
                 plane_band=None, n_obs=1, lr=None, ker_half=0, scene_mode=0, psfs=None, sched=None, prior=None):
        """psfs: dict(obs [S, P, ky, kx], model [1 or S, kym, kxm]) instead of kernel_spec (pass None there): the S * P
        difference kernels are matched on the device and their spectra stay there; ker_idx then indexes s * P + p.
        plane_band / n_obs: several observations per blend -- images, weights and ker_idx then hold the bands of
        all observations stacked as P planes, plane p seeing model band plane_band[p] (include/scarlet_b200.h)."""
        require_device()
        monotonic_tables(Ny, Nx)
        self.C, self.Ny, self.Nx = int(C), int(Ny), int(Nx)
        self.plane_band = None if plane_band is None else np.ascontiguousarray(plane_band, dtype=np.int32)
        # lr: dict(images [B, Plr, n_lr], weights or None, op [n_op, Plr, n_lr, N], op_idx [B]) -- the planes of
        # low-resolution observations; plane_band then lists the full-resolution planes first, these after
        self.lr = lr
        self.psfs = psfs
        # sched: dict(ops int32 [K, 16], args float32 [K, 16]) for components flagged F_SCHEDULE; prior: dict(par [K, 4]
        # = (a_sed, a_morph, L_sed, L_morph; L_sed < 0: none), b_sed [K, C], idx [K], b_morph [n, N]) (scarlet_b200.h)
        self.sched, self.prior = sched, prior
        self.Plr = 0 if lr is None else int(np.shape(lr["images"])[1])
        self.P = self.C if plane_band is None else len(self.plane_band) - self.Plr
        self.n_obs = int(n_obs)
        # scene mode (frames beyond one convolution tile / > 8 components per blend): needs the half-width of the
        # difference kernels for its overlap-save tiles
        self.ker_half, self.scene_mode = int(ker_half), int(scene_mode)
        self.stream = stream
        self._h = c_void_p()
        self._keep = None
        self.max_iter = 0
        desc = self._desc(comp_off, sed, morph, cflags, center, shift, conv, images, weights, kernel_spec, ker_idx,
                          cw, cw_idx, it0, prev_loss, src_first)
        _check(lib().scl_batch_create(ctypes.byref(desc), ctypes.byref(self._h), c_void_p(self.stream)))

    def _desc(self, comp_off, sed, morph, cflags, center, shift, conv, images, weights, kernel_spec, ker_idx,
              cw, cw_idx, it0, prev_loss, src_first=None, keep_spectra=False):
        N = self.Ny * self.Nx
        ca = np.ascontiguousarray
        comp_off = ca(comp_off, dtype=np.int32)
        B = len(comp_off) - 1
        K = int(comp_off[-1]) if B > 0 else 0
        a = dict(
            comp_off=comp_off,
            sed=ca(sed, dtype=np.float32).reshape(K, self.C),
            morph=None if morph is None else ca(morph, dtype=np.float32).reshape(K, N),
            cflags=ca(cflags, dtype=np.uint32).reshape(K),
            center=ca(center, dtype=np.int32).reshape(K, 2),
            shift=ca(shift, dtype=np.float64).reshape(K, 2),
            conv=ca(conv, dtype=np.int32).reshape(K),
            images=ca(images, dtype=np.float32).reshape(B, self.P, N),
            weights=None if weights is None else ca(weights, dtype=np.float32).reshape(B, self.P, N),
            kernel_spec=None if kernel_spec is None else ca(kernel_spec, dtype=np.complex64),
            ker_idx=ca(ker_idx, dtype=np.int32).reshape(B * self.P),
            cw=None if cw is None else ca(cw, dtype=np.float32),
            cw_idx=ca(cw_idx, dtype=np.int32).reshape(B),
            it0=ca(np.zeros(B) if it0 is None else it0, dtype=np.int32).reshape(B),
            prev_loss=None if prev_loss is None else ca(prev_loss, dtype=np.float64).reshape(B),
            src_first=None if src_first is None else ca(src_first, dtype=np.int32).reshape(K),
        )
        self._keep = a
        self.B, self.K = B, K
        d = BatchDesc()
        d.n_blends, d.n_comp, d.C, d.Ny, d.Nx = B, K, self.C, self.Ny, self.Nx
        for name in ("comp_off", "sed", "morph", "cflags", "center", "shift", "conv", "images", "weights",
                     "kernel_spec", "ker_idx", "cw", "cw_idx", "it0", "prev_loss", "src_first"):
            setattr(d, name, _p(a[name]))
        d.ker_half, d.scene_mode = self.ker_half, self.scene_mode
        if self.plane_band is not None:
            d.n_planes, d.plane_band, d.n_obs = self.P, _p(self.plane_band), self.n_obs
        if self.lr is not None:
            lr = dict(images=ca(self.lr["images"], dtype=np.float32),
                      weights=None if self.lr.get("weights") is None else ca(self.lr["weights"], dtype=np.float32),
                      op=ca(self.lr["op"], dtype=np.float32), op_idx=ca(self.lr["op_idx"], dtype=np.int32))
            assert lr["images"].shape[:2] == (B, self.Plr) and lr["op"].shape[1:] == (self.Plr, lr["images"].shape[2], N)
            a.update({"lr_" + k: v for k, v in lr.items()})
            d.n_lr_planes, d.lr_n, d.n_lr_op = self.Plr, lr["images"].shape[2], len(lr["op"])
            d.lr_images, d.lr_weights, d.lr_op, d.lr_op_idx = (_p(lr["images"]), _p(lr["weights"]), _p(lr["op"]),
                                                               _p(lr["op_idx"]))
        d.n_ker = 0 if a["kernel_spec"] is None else len(a["kernel_spec"])
        if self.psfs is not None:
            if a["kernel_spec"] is not None:
                raise ValueError("give kernel_spec or psfs, not both")
            po, pm = np.shape(self.psfs["obs"]), np.shape(self.psfs["model"])
            if len(po) != 4 or len(pm) != 3 or po[1] != self.P:
                raise ValueError("psfs: obs [S, P, ky, kx] and model [1 or S, kym, kxm] expected")
            d.n_ker = po[0] * self.P
            d.ker_half = self.ker_half = self.ker_half or kernel_half(po[2], po[3], self.Ny, self.Nx)
            if not keep_spectra:            # (re-)match on the device; otherwise the batch keeps what it holds
                obs, mod = ca(self.psfs["obs"], dtype=np.float32), ca(self.psfs["model"], dtype=np.float32)
                a.update(psf_obs=obs, psf_model=mod)
                d.psf_obs, d.psf_model = _p(obs), _p(mod)
                d.psf_sets, d.psf_model_sets = po[0], pm[0]
                d.psf_ky, d.psf_kx, d.psf_kym, d.psf_kxm = po[2], po[3], pm[1], pm[2]
        if self.sched is not None:
            a["sched_ops"] = ca(self.sched["ops"], dtype=np.int32).reshape(K, SCHED_OPS)
            a["sched_args"] = ca(self.sched["args"], dtype=np.float32).reshape(K, SCHED_OPS)
            d.sched_ops, d.sched_args = _p(a["sched_ops"]), _p(a["sched_args"])
        if self.prior is not None:
            a["prior_par"] = ca(self.prior["par"], dtype=np.float32).reshape(K, 4)
            a["prior_b_sed"] = ca(self.prior["b_sed"], dtype=np.float32).reshape(K, self.C)
            a["prior_idx"] = ca(self.prior["idx"], dtype=np.int32).reshape(K)
            bm = self.prior.get("b_morph")
            a["prior_b_morph"] = None if bm is None or len(bm) == 0 else ca(bm, dtype=np.float32).reshape(-1, N)
            d.prior_par, d.prior_b_sed, d.prior_idx = _p(a["prior_par"]), _p(a["prior_b_sed"]), _p(a["prior_idx"])
            d.prior_b_morph = _p(a["prior_b_morph"])
            d.n_prior_morph = 0 if a["prior_b_morph"] is None else len(a["prior_b_morph"])
        if a["cw"] is None:
            d.n_cw, d.ph, d.pw = 0, 0, 0
        else:
            d.n_cw, d.ph, d.pw = a["cw"].shape
        self.h2d_bytes = sum(v.nbytes for v in a.values() if v is not None)
        return d

    def upload(self, comp_off, sed, morph, cflags, center, shift, conv, images, weights, kernel_spec, ker_idx,
               cw, cw_idx, it0=None, prev_loss=None, src_first=None, psfs=None):
        """psfs: the PSF sets of the batch; the same object as at construction (or None) keeps the spectra already
        matched on the device, a new one is matched again."""
        keep = self.psfs is not None and (psfs is None or psfs is self.psfs)
        if psfs is not None:
            self.psfs = psfs
        desc = self._desc(comp_off, sed, morph, cflags, center, shift, conv, images, weights, kernel_spec, ker_idx,
                          cw, cw_idx, it0, prev_loss, src_first, keep_spectra=keep)
        _check(lib().scl_batch_upload(self._h, ctypes.byref(desc), c_void_p(self.stream)))

    def fit(self, max_iter=200, e_rel=1e-2, approximate_L=False, check_every=8, profile=False):
        """Returns dict(device_ms, kernel_ms[4], steps, launches)."""
        ms, k1 = c_float(0), (c_float * 4)()
        nk1, nl = c_int(0), c_int(0)
        _check(lib().scl_batch_fit(self._h, int(max_iter), c_double(e_rel), int(bool(approximate_L)), int(check_every),
                                   int(bool(profile)), ctypes.byref(ms), k1, ctypes.byref(nk1),
                                   ctypes.byref(nl), c_void_p(self.stream)))
        self.max_iter = int(max_iter)
        return dict(device_ms=ms.value, kernel_ms=[float(v) for v in k1], steps=nk1.value, launches=nl.value)

    def init_extended(self, cw, jac, cutoff, symmetric=True, monotonic=True):
        """init_extended_source on the device for every component (see include/scarlet_b200.h); returns ok [K] bool."""
        ca = np.ascontiguousarray
        cw = ca(cw, dtype=np.float32).reshape(self.K, self.C)
        jac = ca(jac, dtype=np.float32).reshape(self.K)
        cutoff = ca(cutoff, dtype=np.float32).reshape(self.K)
        ok = np.zeros(self.K, np.int32)
        _check(lib().scl_batch_init_extended(self._h, _p(cw), _p(jac), _p(cutoff), int(bool(symmetric)),
                                             int(bool(monotonic)), _p(ok), c_void_p(self.stream)))
        return ok.astype(bool)

    def init_sources(self, kind, cw, jac, cutoff, symmetric=True, monotonic=True, percentiles=(25,)):
        """scl_batch_init_sources: device initialisation of extended (0), multi-component (1) and point (2) sources,
        kind int32 [K] per component or None (all extended); returns ok [K] bool."""
        ca = np.ascontiguousarray
        cw = ca(cw, dtype=np.float32).reshape(self.K, self.C)
        jac = ca(jac, dtype=np.float32).reshape(self.K)
        cutoff = ca(cutoff, dtype=np.float32).reshape(self.K)
        kind = None if kind is None else ca(kind, dtype=np.int32).reshape(self.K)
        perc = ca(np.sort(np.asarray(percentiles, dtype=np.float32)))
        ok = np.zeros(self.K, np.int32)
        _check(lib().scl_batch_init_sources(self._h, _p(kind), _p(cw), _p(jac), _p(cutoff), int(bool(symmetric)),
                                            int(bool(monotonic)), _p(perc), len(perc), _p(ok), c_void_p(self.stream)))
        return ok.astype(bool)

    def constrain(self):
        """One `update()` pass with it = 0 on every component (what sources run at construction)."""
        _check(lib().scl_batch_constrain(self._h, c_void_p(self.stream)))

    def phase_clocks(self, enable=True):
        """Enable/disable/read the constraint kernel's per-phase clock sums (profiling)."""
        out = (ctypes.c_ulonglong * 8)()
        _check(lib().scl_batch_phase_clocks(self._h, int(bool(enable)), out))
        return [int(v) for v in out]

    def wait_mode(self, blocking):
        """Host waits sleep on a blocking-sync event (True) instead of spinning on the stream (False, default)."""
        _check(lib().scl_batch_wait_mode(self._h, int(bool(blocking))))

    def debug_flags(self, flags):
        """Test switches (include/scarlet_b200.h: scl_batch_debug_flags)."""
        _check(lib().scl_batch_debug_flags(self._h, int(flags)))

    def snapshot(self):
        _check(lib().scl_batch_snapshot(self._h, c_void_p(self.stream)))

    def restore(self):
        _check(lib().scl_batch_restore(self._h, c_void_p(self.stream)))

    def alloc_results(self, max_iter, morph=True):
        """Host arrays `download` fills (allocate once and pass as `out=` to reuse, e.g. page-locked)."""
        return dict(
            sed=np.empty((self.K, self.C), np.float32),
            morph=np.empty((self.K, self.Ny, self.Nx), np.float32) if morph else None,
            center=np.empty((self.K, 2), np.int32),
            shift=np.empty((self.K, 2), np.float64),
            conv=np.empty(self.K, np.int32),
            n_iter=np.empty(self.B, np.int32),
            active=np.empty(self.B, np.int32),
            logL=np.empty((self.B, max_iter), np.float64) if max_iter > 0 else None,
            lipschitz=np.empty((self.B, 2), np.float32),
        )

    def download(self, morph=True, out=None):
        if out is None:
            out = self.alloc_results(self.max_iter, morph)
        elif out["logL"] is not None and out["logL"].shape != (self.B, self.max_iter):
            raise ValueError("out['logL'] must be [B, max_iter of the last fit]")
        _check(lib().scl_batch_download(self._h, _p(out["sed"]), _p(out["morph"]), _p(out["center"]), _p(out["shift"]),
                                        _p(out["conv"]), _p(out["n_iter"]), _p(out["active"]), _p(out["logL"]),
                                        _p(out["lipschitz"]), c_void_p(self.stream)))
        self.d2h_bytes = sum(v.nbytes for v in out.values() if v is not None)
        return out

    def products(self, min_value=0.0):
        """Per component, from the device state: flux [K, C] (Component.get_flux), bounds [K, 4] (bbox.trim of the
        morphology; -1 = nothing above min_value), flags [K] (BlendFlag bits incl. EDGE_PIXELS = 4, NO_VALID_PIXELS = 8)."""
        flux = np.empty((self.K, self.C), np.float32)
        bounds = np.empty((self.K, 4), np.int32)
        flags = np.empty(self.K, np.int32)
        _check(lib().scl_batch_products(self._h, c_float(min_value), _p(flux), _p(bounds), _p(flags), c_void_p(self.stream)))
        return flux, bounds, flags

    def render_components(self, k0=0, nk=None):
        """Observation.render(component.get_model()) of components [k0, k0 + nk): float32 [nk, P, Ny, Nx]."""
        nk = self.K - k0 if nk is None else int(nk)
        out = np.empty((nk, self.P, self.Ny, self.Nx), np.float32)
        _check(lib().scl_batch_render_components(self._h, int(k0), nk, _p(out), c_void_p(self.stream)))
        return out

    def close(self):
        if self._h:
            lib().scl_batch_destroy(self._h)
            self._h = c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
