"""CPU-side checks of the native boundary: the library loads, exports every symbol
include/scarlet_b200.h declares, and the host tables match the oracle."""
import os
import re
import ctypes
import numpy as np
import pytest
from numpy.testing import assert_array_equal

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
