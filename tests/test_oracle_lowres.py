"""The oracle's multi-resolution observation (oracle/ref_lowres.py) against outputs of the reference itself
(tests/golden/make_golden_lowres.py -> ref_golden_lowres.npz): match products, render, loss, adjoint, source
initialisation and a two-observation fit.  CPU only."""
import os
import sys
import numpy as np
import pytest
from numpy.testing import assert_array_equal, assert_allclose

import helpers
from oracle import ref_blend as rb, ref_lowres as rl


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(helpers.GOLD, "ref_golden_lowres.npz"))


@pytest.mark.parametrize("tag", ["A_", "B_", "R_"])
def test_match_render_loss(gold, tag):
    """R_: the low-resolution frame rotated by 10 degrees (observation.py:443-454,483-503)."""
    frame, ohr, olr = helpers.lowres_scene(gold, tag)
    assert list(olr._fft_shape) == list(gold[tag + "fft_shape"])
    assert tuple(olr.lr_shape) == tuple(gold[tag + "lr_shape"])
    assert bool(olr.small_axis) == bool(gold[tag + "small_axis"])
    assert_array_equal(olr._coord_lr[0], gold[tag + "coord_lr0"])             # index work: exact
    assert_array_equal(olr._coord_lr[1], gold[tag + "coord_lr1"])
    assert_allclose(olr._coord_hr[0], gold[tag + "coord_hr0"], rtol=0, atol=1e-12)
    assert_allclose(np.array(olr.other_shifts), gold[tag + "other_shifts"], rtol=0, atol=1e-12)
    ref = gold[tag + "resconv_op"]
    assert_allclose(olr._resconv_op, ref, rtol=0, atol=2e-6 * np.abs(ref).max())
    model = gold[tag + "model"]
    ref = gold[tag + "render"]
    assert_allclose(olr._render(model), ref, rtol=0, atol=1e-5 * np.abs(ref).max())
    assert_allclose(olr.get_loss(model), gold[tag + "loss"], rtol=1e-5)


def test_diffkernel_and_adjoint(gold):
    frame, ohr, olr = helpers.lowres_scene(gold, "A_")
    ref = gold["A_diff_psf"]
    assert_allclose(olr.build_diffkernel(frame), ref, rtol=0, atol=2e-6 * np.abs(ref).max())
    g = olr.grad(gold["A_model"])
    assert not g[:3].any()
    ref = gold["A_grad"]
    assert_allclose(g[3:], ref, rtol=0, atol=2e-5 * np.abs(ref).max())
    # finite differences of the restated loss, float64
    m = gold["A_model"].astype(np.float64)
    frame64, _, o64 = helpers.lowres_scene(gold, "A_", dtype=np.float64)
    g64 = o64.grad(m)
    rng = np.random.default_rng(0)
    d = rng.normal(size=m.shape)
    eps = 1e-4
    fd = (o64.get_loss(m + eps * d) - o64.get_loss(m - eps * d)) / (2 * eps)
    assert_allclose(np.sum(g64 * d), fd, rtol=1e-7)


def test_combined_source_and_fit(gold):
    frame, ohr, olr = helpers.lowres_scene(gold, "A_")
    srcs, blend = helpers.lowres_blend(gold, frame, ohr, olr)
    assert_array_equal([s.pixel_center for s in srcs], gold["A_init_centers"])
    assert_allclose(np.array([c.sed for c in blend.components]), gold["A_init_seds"], rtol=2e-6)
    assert_allclose(np.array([c.morph for c in blend.components]), gold["A_init_morphs"], rtol=0, atol=2e-6)
    ref = gold["A_mse"]
    for _ in range(len(ref)):
        blend.step(1e-9)
    assert_allclose(blend.mse, ref, rtol=1e-4)
    assert_array_equal([s.pixel_center for s in srcs], gold["A_centers"])
    gm = gold["A_fit_model"]
    assert_allclose(blend.get_model(), gm, rtol=0, atol=1e-4 * np.abs(gm).max())


def test_rotated_diffkernel_adjoint_and_product_operator(gold):
    """Rotated frames: the difference kernel through the rotated sinc interpolation (interpolation.py:374-404) and the
    adjoint against the reference's own; and the matrix the product's LowResObservation.match builds for the device
    (one FFT product per band, row and column, scarlet_b200/observation.py:_match_rotated) against the oracle's
    `_render` applied to every unit model."""
    frame, ohr, olr = helpers.lowres_scene(gold, "R_")
    assert olr.isrot
    ref = gold["R_diff_psf"]
    assert_allclose(olr.build_diffkernel(frame, (olr.cos_rot, olr.sin_rot)), ref, rtol=0, atol=2e-6 * np.abs(ref).max())
    g = olr.grad(gold["R_model"])
    ref = gold["R_grad"]
    assert_allclose(g[3:], ref, rtol=0, atol=2e-5 * np.abs(ref).max())
    import scarlet
    helpers.sys_path_golden()
    from wcs_stub import make_pair
    hr_img, lr_img = gold["R_hr_img"], gold["R_lr_img"]
    whr, wlr = make_pair(hr_img.shape[1:], lr_img.shape[1:], scale=4.0, angle_deg=float(gold["R_angle"]),
                         lr_offset=tuple(gold["R_offset"]))
    chan = ["h0", "h1", "h2", "l0", "l1", "l2", "l3", "l4"]
    pf = scarlet.Frame((8,) + hr_img.shape[1:], wcs=whr, psfs=gold["R_mpsf"][None], channels=chan)
    po = scarlet.LowResObservation(lr_img, wcs=wlr, psfs=gold["R_lr_psfs"], weights=gold["R_w_lr"],
                                   channels=chan[3:]).match(pf)
    assert po.isrot and po.lr_shape == tuple(int(v) for v in gold["R_lr_shape"])
    L = olr.dense_operator(frame.shape)
    assert_allclose(po._lr_op, L, rtol=0, atol=1e-6 * np.abs(L).max())
    ref = gold["R_resconv_op"]
    assert_allclose(po._resconv_op, ref, rtol=0, atol=2e-6 * np.abs(ref).max())
