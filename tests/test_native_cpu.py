"""CPU-side checks of the native boundary: the library loads, exports every symbol
include/scarlet_b200.h declares, and the host tables match the oracle."""
import os
import re
import ctypes
import numpy as np
import pytest
from numpy.testing import assert_array_equal, assert_allclose

from oracle import ref_prox, ref_fft

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_header_symbols():
    from scarlet_b200 import _native, build
    build.build()
    L = _native.lib()
    hdr = open(os.path.join(ROOT, "include", "scarlet_b200.h")).read()
    names = set(re.findall(r"\b(scl_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) >= 15
    for n in names:
        assert hasattr(L, n), n
    assert L.scl_version() == 100


def test_no_device_is_loud():
    from scarlet_b200 import _native
    if _native.lib().scl_device_count() > 0:
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError):
        _native.prox_weighted_monotonic(np.zeros((1, 5, 5), np.float32), [(2, 2)])


@pytest.mark.parametrize("shape,center", [((9, 7), (2, 5)), ((16, 13), (3, 3)), ((12, 12), (0, 11)), ((5, 5), (2, 2))])
def test_offset_tables_reproduce_reference_weights(shape, center):
    """wt9/ctab (offset-indexed) reproduce getRadialMonotonicWeights bit for bit when evaluated the way the
    kernels do (f32 path: wt9[border type]; f64 path: ctab renormalised in tap order)."""
    from scarlet_b200 import _native
    wt9, ctab = _native.monotonic_tables_host(*shape)
    H, W = shape
    ref = ref_prox.monotonic_weights(shape, center)
    for y in range(H):
        for x in range(W):
            Y, X = y - center[0] + H - 1, x - center[1] + W - 1
            by = 1 if y == 0 else (2 if y == H - 1 else 0)
            bx = 1 if x == 0 else (2 if x == W - 1 else 0)
            assert_array_equal(wt9[by * 3 + bx, Y, X], ref[:, y * W + x].astype(np.float32))
            c = ctab[Y, X].copy()
            for i, (oy, ox) in enumerate(ref_prox.COORDS):
                if not (0 <= y + oy < H and 0 <= x + ox < W):
                    c[i] = 0
            s = 0.0
            for i in range(8):
                s = s + c[i]
            assert_array_equal(c / (s if s != 0 else 1), ref[:, y * W + x])


def test_wavefront_levels_are_a_topological_order():
    """level = 2 max(|Y|,|X|) + min(|Y|,|X|) - 1: every non-zero weight points to a strictly lower level, and
    sweeping level by level (random order inside) equals the radius-sorted sequential sweep bit for bit."""
    rng = np.random.default_rng(0)
    for shape, center in [((64, 64), (20, 41)), ((17, 30), (16, 0)), ((40, 59), (5, 58))]:
        H, W = shape
        Y, X = np.meshgrid(np.arange(H) - center[0], np.arange(W) - center[1], indexing="ij")
        lv = (2 * np.maximum(abs(Y), abs(X)) + np.minimum(abs(Y), abs(X)) - 1).ravel()
        w = ref_prox.monotonic_weights(shape, center)
        offs = ref_prox.offsets(W)
        for i in range(8):
            p = np.nonzero(w[i] > 0)[0]
            assert np.all(lv[p + offs[i]] < lv[p])
        x = rng.random(shape).astype(np.float32)
        a = x.copy()
        ref_prox.build_monotonic(shape, center, .1)(a)
        order = np.lexsort((rng.random(lv.size), lv))
        order = order[lv[order] >= 0]
        b = x.copy()
        ref_prox.prox_weighted_monotonic(b, w, order, offs, .1)
        assert_array_equal(a, b)


def test_kernel_spectrum_layout():
    from scarlet_b200 import _native
    rng = np.random.default_rng(1)
    ker = rng.random((2, 7, 7))
    spec = _native.kernel_spectrum(ker, 20, 30)
    assert spec.shape == (2, 128, 65) and spec.dtype == np.complex64
    img = rng.random((20, 30))
    pad = np.zeros((128, 128))
    pad[:20, :30] = img
    got = np.fft.irfft2(np.fft.rfft2(pad) * spec[1].astype(np.complex128), s=(128, 128))[:20, :30]
    ref = ref_fft.convolve(img[None], ker[1:2], axes=(1, 2))[0]
    np.testing.assert_allclose(got, ref, atol=1e-5)
    assert [_native.next_fast_len(n) for n in range(1, 300)] == [ref_fft.next_fast_len(n) for n in range(1, 300)]


def test_box_container():
    """reference tests/test_bbox.py:6-95: construction, bounds, slices, empty boxes, copy, ==, &, |."""
    from scarlet_b200.bbox import Box
    b1 = Box((4, 1), 7, 6)
    assert (b1.height, b1.width, b1.bottom, b1.top, b1.left, b1.right) == (7, 6, 4, 10, 1, 6)
    assert b1.slices == (slice(4, 11), slice(1, 7)) and b1.shape == (7, 6) and b1.is_empty is False
    b2 = Box.from_bounds(4, 10, 1, 6)
    assert b2 == b1 and b2.slices == b1.slices and b2.shape == b1.shape
    e = Box.from_bounds(10, 9, 5, 10)
    assert e.is_empty is True and (e.width, e.height) == (0, 0)
    assert e.bottom is None and e.top is None and e.left is None and e.right is None and e.yx0 is None
    assert e.slices == (slice(0, 0), slice(0, 0))

    b1 = Box((5, 10), 9, 7)
    c = b1.copy()
    c.yx0, c.height, c.width = (0, 2), 1, 3
    assert (b1.bottom, b1.top, b1.left, b1.right) == (c.bottom + 5, c.top + 13, c.left + 8, c.right + 12)
    c = b1.copy()
    assert b1 == c and (b1 & c) == b1 and (b1 | c) == b1
    lo = Box((3, 7), 9, 8)
    assert b1 != lo
    assert (b1 & lo) == Box.from_bounds(b1.bottom, lo.top, b1.left, lo.right)
    assert (b1 | lo) == Box.from_bounds(lo.bottom, b1.top, lo.left, b1.right)
    hi = Box((7, 12), 11, 13)
    assert (b1 & hi) == Box.from_bounds(hi.bottom, b1.top, hi.left, b1.right)
    assert (b1 | hi) == Box.from_bounds(b1.bottom, hi.top, b1.left, hi.right)
    far = Box((30, 50), 10, 11)
    assert (b1 & far) == Box((0, 0), width=0, height=0) and (b1 & far).is_empty is True
    assert (b1 | far) == Box.from_bounds(b1.bottom, far.top, b1.left, far.right)


def test_interpolation_taps_vs_reference_fixture():
    """The 1-D interpolation kernels (host parameters of scl_resample_f32) against the reference's own taps and
    windows (tests/golden/make_golden_resample.py; interpolation.py:139-270)."""
    from scarlet_b200 import interpolation
    g = np.load(os.path.join(ROOT, "tests", "golden", "ref_golden_resample.npz"))
    for name in ("lanczos", "bilinear", "catmull_rom", "mitchel_netravali", "cubic_spline", "quintic_spline"):
        fn = getattr(interpolation, name)
        for taps, d in zip(g["taps_" + name], (.217, -.691, .5, 0.)):
            y, w = fn(d)
            assert_allclose(np.concatenate([y, w]), taps, rtol=0, atol=1e-14)
            assert abs(y.sum() - 1) < 1e-2 or name == "lanczos"
    k, yw, xw = interpolation.get_separable_kernel(.3, -.2, kernel=interpolation.bilinear)
    assert_allclose(k, np.outer([.7, .3], [.2, .8]))
    assert list(yw) == [0, 1] and list(xw) == [-1, 0]
    with pytest.raises(ValueError):
        interpolation.bilinear(-1.01)
