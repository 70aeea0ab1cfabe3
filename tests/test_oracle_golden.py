"""The oracle against (a) the reference's own known-answer vectors and (b)
outputs of the reference run in the build container (tests/golden)."""
import os
import numpy as np
import pytest
from numpy.testing import assert_array_equal, assert_almost_equal, assert_allclose

from oracle import ref_fft, ref_prox, ref_blend

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ---------------------------------------------------------------- reference KATs
def test_kat_prox_monotonic():
    """reference tests/test_operator.py:10-56"""
    X = np.arange(25).reshape(5, 5).astype(np.float64)
    ref_didx = [12, 17, 13, 7, 11, 18, 16, 6, 8, 14, 10, 22, 2, 9, 23, 5, 15, 3, 1, 19, 21, 0, 4, 20, 24]
    r2 = ref_prox.radius2((5, 5), (2, 2))
    didx = ref_prox.sort_by_radius((5, 5), None, kind="stable")
    # identical r^2 level sequence; tie order inside a level is numpy-version dependent
    assert_array_equal(r2[didx], r2[ref_didx])
    assert ref_prox.nearest_ref_idx((5, 5)).tolist() == [6, 7, 7, 7, 8, 11, 12, 12, 12, 13, 11, 12, 12,
                                                         12, 13, 11, 12, 12, 12, 13, 16, 17, 17, 17, 18]
    assert_array_equal(ref_prox.offsets(5), [-6, -5, -4, -1, 1, 4, 5, 6])
    Y = X.copy()
    ref_prox.prox_nearest_monotonic(Y, ref_prox.nearest_ref_idx((5, 5)), ref_didx)
    assert_array_equal(Y, [[0, 1, 2, 3, 4], [5, 6, 7, 8, 9], [10, 11, 12, 12, 12],
                           [11, 12, 12, 12, 12], [12, 12, 12, 12, 12]])
    Y = X.copy()
    ref_prox.build_monotonic((5, 5), None, 0)(Y)
    assert_almost_equal(Y, [[0., 1., 2., 3., 4.], [5., 6., 7., 8., 9.],
                            [9.74264069, 11., 12., 12., 10.82842712],
                            [11.0306277, 11.70710678, 12., 12., 11.77123617],
                            [11.55634919, 11.86886724, 11.91421356, 11.98324916, 11.92809042]])
    Y = X.copy()
    ref_prox.build_monotonic((5, 5), None, .25)(Y)
    assert_almost_equal(Y, [[0.000000000, 1.000000000, 2.000000000, 3.000000000, 4.000000000],
                            [5.000000000, 6.000000000, 7.000000000, 7.242640687, 5.806841831],
                            [5.801461031, 9.000000000, 12.000000000, 9.000000000, 6.074431804],
                            [5.895545844, 7.681980515, 9.000000000, 7.681980515, 5.935521488],
                            [4.988519641, 5.949655012, 6.170941546, 5.949655012, 4.997301087]])
    # order independence inside r^2 ties: the reference's pinned order gives the same bits
    Z = X.copy()
    ref_prox.prox_weighted_monotonic(Z, ref_prox.monotonic_weights((5, 5)), ref_didx[1:], ref_prox.offsets(5), .25)
    assert_array_equal(Y, Z)


def test_kat_symmetry():
    """reference tests/test_operator.py:92-123, tests/test_update.py:204-211"""
    X = np.arange(25, dtype=float).reshape(5, 5)
    Y = X.copy()
    ref_prox.prox_soft_symmetry(Y, 0, strength=1)
    assert_array_equal(Y, np.full((5, 5), 12.0))
    X = np.zeros((21, 21))
    X[8:13, 8:13] = [[1, 2, 3, 2, 1], [2, 3, 4, 3, 1], [3, 4, 5, 1, 1], [2, 3, 1, 1, 1], [1, 1, 1, 1, 1]]
    Y = ref_prox.prox_kspace_symmetry(X.copy(), 0, shift=(0, 0))
    assert_almost_equal(Y, (X + X[::-1, ::-1]) / 2)
    Y = X.copy()
    ref_prox.prox_soft_symmetry(Y, 0, .5)
    assert_almost_equal(Y, .75 * X + .25 * X[::-1, ::-1])
    X = np.arange(25, dtype=float).reshape(5, 5)
    ref_prox.prox_uncentered_symmetry(X, 0, (1, 1), "soft", None, None, 1)
    assert_array_equal(X[:3, :3], np.full((3, 3), 6.0))
    assert_array_equal(X[3:], np.arange(25).reshape(5, 5)[3:])


def test_kat_uncentered_window():
    """reference tests/test_operator.py:131-219 (window geometry)"""
    assert ref_prox.uncentered_slices((5, 5), (2, 2)) is None
    assert ref_prox.uncentered_slices((5, 5), (1, 1)) == (slice(None, -2), slice(None, -2))
    assert ref_prox.uncentered_slices((5, 5), (3, 3)) == (slice(2, None), slice(2, None))
    assert ref_prox.uncentered_slices((6, 6), (2, 2)) == (slice(None, -1), slice(None, -1))
    assert ref_prox.uncentered_slices((6, 6), (4, 1)) == (slice(3, None), slice(None, -3))


def test_kat_pad_centered():
    """reference tests/test_fft.py:9-45: odd array in even box lands centre-right"""
    x = np.arange(9.).reshape(3, 3)
    p = ref_fft.pad(x, (6, 6))
    assert p[3, 3] == 4 and p.sum() == x.sum()
    assert_array_equal(ref_fft.centered(p, (3, 3)), x)
    p = ref_fft.pad(x, (4, 7))
    assert p[2, 3] == 4
    assert [ref_fft.next_fast_len(n) for n in (85, 90, 97, 109, 128, 138, 7)] == [90, 90, 100, 120, 128, 144, 8]
    assert ref_fft.fft_shape((5, 69, 57), (5, 41, 41), 3, (1, 2)) == [120, 108]
    assert ref_fft.fft_shape((6, 40, 59), (6, 31, 31), 3, (1, 2)) == [75, 96]


def test_kat_max_pixel_plus_normalized():
    """reference tests/test_update.py:11-68"""
    m = np.zeros((9, 9))
    m[4, 4] = 1
    m[2, 5] = 2
    m[0, 0] = 10
    assert ref_prox.max_pixel(m, (4, 4)) == (2, 5)
    sed = np.array([1., -2., 3.])
    morph = np.array([[-1., 2.], [4., -3.]])
    ref_prox.prox_plus(sed)
    ref_prox.prox_plus(morph)
    assert_array_equal(sed, [1, 0, 3])
    ref_prox.normalized(sed, morph, "morph_max")
    assert_array_equal(morph, [[0, .5], [1, 0]])
    assert_array_equal(sed, [4, 0, 12])


# ---------------------------------------------------------------- reference outputs
@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_monotonic_vs_reference(golden, tag):
    shape = tuple(golden["mono_%s_shape" % tag])
    center = tuple(golden["mono_%s_center" % tag])
    w = ref_prox.monotonic_weights(shape, center)
    assert_array_equal(w, golden["mono_%s_weights" % tag])            # bit-exact f64
    r2 = ref_prox.radius2(shape, center)
    assert_array_equal(r2[ref_prox.sort_by_radius(shape, center, kind="stable")],
                       r2[golden["mono_%s_didx" % tag]])              # level sets
    for dt in ("float32", "float64"):
        for th, t in ((0, 0), (.1, 1)):
            x = golden["mono_%s_in_%s" % (tag, dt)].copy()
            ref_prox.build_monotonic(shape, center, th)(x)
            assert_array_equal(x, golden["mono_%s_out_%s_t%d" % (tag, dt, t)])  # bit-exact
    x = golden["mono_%s_in_float32" % tag].copy()
    y = x.copy()
    ref_prox.prox_weighted_monotonic_py(y, w, ref_prox.sort_by_radius(shape, center, "stable")[1:],
                                        ref_prox.offsets(shape[1]), .1)
    assert_array_equal(y, golden["mono_%s_out_float32_t1" % tag])


@pytest.mark.parametrize("tag", ["a", "b", "c", "d", "e"])
def test_symmetry_vs_reference(golden, tag):
    x = golden["sym_%s_in" % tag]
    c = tuple(golden["sym_%s_center" % tag])
    shift = tuple(golden["sym_%s_shift" % tag])
    y = x.copy()
    ref_prox.prox_uncentered_symmetry(y, 0, c, "kspace", None, shift)
    assert_array_equal(y, golden["sym_%s_kspace" % tag])
    y = x.copy()
    ref_prox.prox_uncentered_symmetry(y, 0, c, "soft", None, None, 1)
    assert_array_equal(y, golden["sym_%s_soft" % tag])
    y = x.copy()
    ref_prox.prox_uncentered_symmetry(y, 0, c, "sdss")
    assert_array_equal(y, golden["sym_%s_sdss" % tag])


@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_measurement_vs_reference(golden, tag):
    m = golden["meas_%s_in" % tag]
    c = tuple(golden["meas_%s_center" % tag])
    assert_array_equal(ref_prox.max_pixel(m, c), golden["meas_%s_maxpix" % tag])
    cc, s = ref_prox.psf_weighted_centroid(m, golden["psf_target_41"].astype(np.float32), c)
    assert_array_equal(cc, golden["meas_%s_centroid" % tag])
    assert_array_equal(np.array(s, dtype=np.float64), golden["meas_%s_shift" % tag])


def test_psf_image_vs_reference(golden):
    psf = ref_blend.generate_psf_image(ref_blend.gaussian, (41, 41), sigma=.9)
    assert_array_equal(psf, golden["psf_target_41"])
