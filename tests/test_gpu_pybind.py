"""The same-named stand-in for the reference's native module (scarlet.operators_pybind11,
/root/reference/scarlet/operators_pybind11.cc:72-91) on the device: the bodies of the reference's
tests/test_operator.py:10-56 with the operator tables built the way operator.prox_strict_monotonic builds them
(oracle.ref_prox restates operator.py:40-122,540-621), handed to the native functions with the reference's argument
lists (operator.py:24-37)."""
import numpy as np
import pytest
from numpy.testing import assert_array_equal, assert_almost_equal

from oracle import ref_prox

pytestmark = pytest.mark.gpu

# the pinned (old-numpy) tie order of tests/test_operator.py:14: any order with the same r^2 level sets gives the
# same answer, the native loop just walks what it is given
DIDX = [12, 17, 13, 7, 11, 18, 16, 6, 8, 14, 10, 22, 2, 9, 23, 5, 15, 3, 1, 19, 21, 0, 4, 20, 24]


@pytest.fixture(scope="module")
def pb():
    import scarlet
    from scarlet import operators_pybind11
    scarlet._native.require_device()
    assert scarlet.operators_pybind11 is operators_pybind11
    return operators_pybind11


def test_reference_kat_nearest(pb):
    """tests/test_operator.py:11-27"""
    X = np.arange(25).reshape(5, 5).astype(np.float64)
    ref_idx = list(ref_prox.nearest_ref_idx(X.shape))
    assert ref_idx == [6, 7, 7, 7, 8, 11, 12, 12, 12, 13, 11, 12, 12, 12, 13, 11, 12, 12, 12, 13, 16, 17, 17, 17, 18]
    _X = X.copy()
    assert pb.prox_monotonic(_X.reshape(-1), 0.0, ref_idx, DIDX, 0) is None
    assert_array_equal(_X, [[0, 1, 2, 3, 4], [5, 6, 7, 8, 9], [10, 11, 12, 12, 12], [11, 12, 12, 12, 12],
                            [12, 12, 12, 12, 12]])


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_reference_kat_weighted(pb, dtype):
    """tests/test_operator.py:33-56 (the float32 overload against the same goldens at float32 precision)"""
    X = np.arange(25).reshape(5, 5).astype(dtype)
    weights = ref_prox.monotonic_weights(X.shape)
    offsets = ref_prox.offsets(5)
    assert_array_equal(offsets, [-6, -5, -4, -1, 1, 4, 5, 6])
    dec = 7 if dtype == np.float64 else 5
    _X = X.copy()
    pb.prox_weighted_monotonic(_X.reshape(-1), 0.0, weights, offsets, DIDX[1:], 0)
    assert_almost_equal(_X, [[0., 1., 2., 3., 4.], [5., 6., 7., 8., 9.],
                             [9.74264069, 11., 12., 12., 10.82842712],
                             [11.0306277, 11.70710678, 12., 12., 11.77123617],
                             [11.55634919, 11.86886724, 11.91421356, 11.98324916, 11.92809042]], decimal=dec)
    _X = X.copy()
    pb.prox_weighted_monotonic(_X.reshape(-1), 0.0, weights, offsets, DIDX[1:], .25)
    assert_almost_equal(_X, [[0.000000000, 1.000000000, 2.000000000, 3.000000000, 4.000000000],
                             [5.000000000, 6.000000000, 7.000000000, 7.242640687, 5.806841831],
                             [5.801461031, 9.000000000, 12.000000000, 9.000000000, 6.074431804],
                             [5.895545844, 7.681980515, 9.000000000, 7.681980515, 5.935521488],
                             [4.988519641, 5.949655012, 6.170941546, 5.949655012, 4.997301087]], decimal=dec)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("shape,center", [((64, 64), (20, 41)), ((69, 57), (3, 55)), ((7, 100), (6, 0))])
def test_caller_tables_equal_oracle_and_device_tables(pb, dtype, shape, center):
    """The sequential walk on caller tables == the oracle's C restatement of the same loop == the wavefront on the
    library's own device-generated tables, bit for bit."""
    from scarlet_b200 import _native as nat
    rng = np.random.default_rng(7)
    X = rng.random(shape).astype(dtype)
    weights = ref_prox.monotonic_weights(shape, center)
    offsets = ref_prox.offsets(shape[1])
    didx = ref_prox.sort_by_radius(shape, center)
    for thresh in (0, .1):
        got = X.copy()
        pb.prox_weighted_monotonic(got.reshape(-1), 0, weights, offsets, didx[1:], thresh)
        want = X.copy()
        ref_prox.prox_weighted_monotonic(want, weights, didx[1:], offsets, thresh)
        assert_array_equal(got, want)
        dev = X.copy()[None]
        nat.prox_weighted_monotonic(dev, [center], thresh)
        assert_array_equal(got, dev[0])


def test_binding_behaviour(pb):
    """pybind11's no-copy rules: wrong dtype / rank is a TypeError; indices that leave the image are refused."""
    w = ref_prox.monotonic_weights((5, 5))
    off = ref_prox.offsets(5)
    with pytest.raises(TypeError):
        pb.prox_weighted_monotonic(np.zeros(25, np.int32), 0, w, off, DIDX[1:], 0)
    with pytest.raises(TypeError):
        pb.prox_weighted_monotonic(np.zeros((5, 5), np.float32), 0, w, off, DIDX[1:], 0)
    with pytest.raises(TypeError):
        pb.prox_monotonic(np.zeros(25, np.float32), 0, list(range(25)), DIDX, 0)
    with pytest.raises(ValueError):
        pb.prox_weighted_monotonic(np.zeros(25), 0, w, off, [25], 0)
    bad = w.copy()
    bad[0, 0] = 1.0                    # tap (-1,-1) of pixel 0 leaves the image
    with pytest.raises(ValueError):
        pb.prox_weighted_monotonic(np.zeros(25), 0, bad, off, [0], 0)
    x = np.arange(25, dtype=np.float64)
    pb.prox_weighted_monotonic(x, 0, w, off, [], 0)          # empty order: untouched
    assert_array_equal(x, np.arange(25))
    strided = np.arange(50, dtype=np.float64)[::2]           # a writeable 1-D view is fine for Eigen::Ref<..., InnerStride>
    keep = strided.copy()
    pb.prox_monotonic(strided, 0, list(ref_prox.nearest_ref_idx((5, 5))), DIDX, 0)
    want = keep.copy()
    ref_prox.prox_nearest_monotonic(want.reshape(5, 5), ref_prox.nearest_ref_idx((5, 5)), DIDX, 0)
    assert_array_equal(strided, want)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_apply_filter(pb, dtype):
    """operators_pybind11.cc:52-70 against its definition in numpy."""
    rng = np.random.default_rng(3)
    img = rng.random((13, 17)).astype(dtype)
    vals = rng.random(5).astype(dtype)
    ys, ye = np.array([0, 1, 0, 2, 1]), np.array([1, 0, 0, 0, 3])
    xs, xe = np.array([1, 0, 0, 0, 2]), np.array([0, 1, 0, 4, 0])
    res = np.full(img.shape, 7, dtype=dtype)
    pb.apply_filter(img, vals, ys, ye, xs, xe, res)
    want = np.zeros(img.shape, dtype)
    for n in range(5):
        r, c = img.shape[0] - ys[n] - ye[n], img.shape[1] - xs[n] - xe[n]
        want[ys[n]:ys[n] + r, xs[n]:xs[n] + c] += vals[n] * img[ye[n]:ye[n] + r, xe[n]:xe[n] + c]
    assert_array_equal(res, want)
