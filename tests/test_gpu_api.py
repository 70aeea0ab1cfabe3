"""The drop-in Python surface (scarlet_b200 / `import scarlet`) on the GPU, written like the
reference's own tests (tests/test_blend.py, test_operator.py, test_update.py, test_source.py)."""
import os
import numpy as np
import pytest
from numpy.testing import assert_array_equal, assert_allclose, assert_almost_equal

import helpers

pytestmark = pytest.mark.gpu
GOLD = helpers.GOLD


@pytest.fixture(scope="module")
def scarlet():
    import scarlet
    scarlet._native.require_device()
    return scarlet


def _scene(scarlet, tag):
    fn = {"hsc": "hsc_cosmos_inputs.npz", "sim": "psf_unmatched_sim_inputs.npz"}[tag]
    d = np.load(os.path.join(GOLD, fn))
    images, psfs = d["images"], d["psfs"]
    mpsf = scarlet.psf.generate_psf_image(scarlet.psf.gaussian, psfs[0].shape, sigma=.9).image
    frame = scarlet.Frame(images.shape, psfs=mpsf[None])
    obs = scarlet.Observation(images, psfs=psfs).match(frame)
    bg = np.ones(len(images), dtype=images.dtype) * float(d["bg_rms"])
    srcs = []
    for (y, x), star in zip(d["coords"], d["is_star"]):
        if star:
            srcs.append(scarlet.PointSource(frame, (y, x), obs, fix_morph=True))
        else:
            srcs.append(scarlet.ExtendedSource(frame, (y, x), obs, bg))
    return frame, obs, srcs, scarlet.Blend(srcs, obs)


@pytest.mark.parametrize("tag", ["hsc", "sim"])
def test_quickstart_workflow(scarlet, golden, tag):
    """docs/quickstart.ipynb cells 3-14 / docs/tutorials/point_source.ipynb cells 2-12 vs the reference's run."""
    frame, obs, srcs, blend = _scene(scarlet, tag)
    assert_allclose(obs._diff_kernels.image, golden[tag + "_diff_kernels"], rtol=0, atol=1e-7)
    seds = np.array([c.sed for c in blend.components])
    morphs = np.array([c.morph for c in blend.components])
    assert_allclose(seds, golden[tag + "_init_seds"], rtol=2e-6)
    assert_allclose(morphs, golden[tag + "_init_morphs"], rtol=0, atol=3e-6)
    assert_array_equal([s.pixel_center for s in srcs], golden[tag + "_init_centers"])
    assert_allclose([s.shift for s in srcs], golden[tag + "_init_shifts"], rtol=0, atol=5e-6)
    blend.fit(200)
    ref = golden[tag + "_mse"]
    assert blend.it == len(ref)
    assert blend.converged
    assert_allclose(blend.mse, ref, rtol=5e-5)
    assert_array_equal([s.pixel_center for s in srcs], golden[tag + "_centers"])
    gm = golden[tag + "_model"]
    assert_allclose(blend.get_model(), gm, rtol=0, atol=5e-5 * np.abs(gm).max())
    # post-fit product: observation.render(blend.get_model()) (tests/test_blend.py:68)
    rendered = obs.render(blend.get_model())
    assert rendered.shape == obs.images.shape
    assert abs(obs.get_loss(blend.get_model()) / blend.mse[-1] - 1) < .2


def test_multicomponent_source_workflow(scarlet):
    """MultiComponentSource (source.py:537-641) through the drop-in surface: construction (init + update at it = 0)
    and Blend.fit, against the reference's own run (tests/golden/make_golden_multi.py)."""
    d = np.load(os.path.join(GOLD, "hsc_cosmos_inputs.npz"))
    g = np.load(os.path.join(GOLD, "ref_golden_multi.npz"))
    images, psfs = d["images"], d["psfs"]
    mpsf = scarlet.psf.generate_psf_image(scarlet.psf.gaussian, psfs[0].shape, sigma=.9).image
    frame = scarlet.Frame(images.shape, psfs=mpsf[None])
    obs = scarlet.Observation(images, psfs=psfs).match(frame)
    bg = np.ones(len(images), dtype=images.dtype) * float(d["bg_rms"])
    srcs = [(scarlet.MultiComponentSource if i < int(g["n_multi"]) else scarlet.ExtendedSource)(frame, (y, x), obs, bg)
            for i, (y, x) in enumerate(d["coords"])]
    assert [getattr(s, "n_components", 1) for s in srcs] == list(g["n_comp"])
    blend = scarlet.Blend(srcs, obs)
    assert_allclose([c.sed for c in blend.components], g["init_seds"], rtol=2e-6)
    assert_allclose([c.morph for c in blend.components], g["init_morphs"], rtol=0, atol=3e-6)
    assert_array_equal([s.pixel_center for s in srcs], g["init_centers"])
    blend.fit(int(g["max_iter"]))
    assert blend.it == len(g["mse"])
    assert blend.converged
    assert_allclose(blend.mse, g["mse"], rtol=5e-5)
    assert_array_equal([s.pixel_center for s in srcs], g["centers"])
    assert_allclose(blend.get_model(), g["model"], rtol=0, atol=5e-5 * np.abs(g["model"]).max())


def test_fit_again_continues(scarlet):
    frame, obs, srcs, blend = _scene(scarlet, "hsc")
    blend.fit(3)
    assert blend.it == 3
    blend.fit(200)
    frame, obs, srcs, blend2 = _scene(scarlet, "hsc")
    blend2.fit(200)
    assert blend.it == blend2.it
    assert_allclose(blend.mse, blend2.mse, rtol=1e-6)


def test_extended_sources_batched_equals_single(scarlet):
    frame, obs, srcs, blend = _scene(scarlet, "hsc")
    d = np.load(os.path.join(GOLD, "hsc_cosmos_inputs.npz"))
    bg = np.ones(5, dtype=np.float32) * .1
    many = scarlet.extended_sources(frame, [tuple(c) for c in d["coords"]], obs, bg)
    assert len(many) == len(srcs)
    for a, b in zip(many, srcs):
        assert_array_equal(a.morph, b.morph)
        assert_array_equal(a.sed, b.sed)
        assert a.pixel_center == b.pixel_center and a.shift == b.shift


def test_fit_many_equals_fit(scarlet):
    blends = [_scene(scarlet, "hsc")[3] for _ in range(3)]
    blends[1].components[0]._sed *= 1.5
    scarlet.fit_many(blends, 200)
    single = _scene(scarlet, "hsc")[3]
    single.fit(200)
    assert_array_equal(blends[0].mse, single.mse)
    assert_array_equal(blends[2].components[3].morph, single.components[3].morph)
    assert blends[1].mse[0] != single.mse[0]


def test_operator_surface(scarlet):
    """reference tests/test_operator.py:10-123, tests/test_update.py:11-68,204-211 through the device"""
    X = np.arange(25).reshape(5, 5).astype(np.float64)
    prox = scarlet.operator.prox_strict_monotonic(X.shape, use_nearest=True, thresh=0)
    _X = X.copy()
    assert prox(_X, 0.0) is _X
    assert_array_equal(_X, [[0, 1, 2, 3, 4], [5, 6, 7, 8, 9], [10, 11, 12, 12, 12],
                            [11, 12, 12, 12, 12], [12, 12, 12, 12, 12]])
    with pytest.raises(ValueError):
        scarlet.operator.prox_strict_monotonic(X.shape, use_nearest=True, thresh=.25)
    prox = scarlet.operator.prox_strict_monotonic(X.shape, use_nearest=False, thresh=.25)
    _X = X.copy()
    prox(_X, 0.0)
    assert_almost_equal(_X, [[0.000000000, 1.000000000, 2.000000000, 3.000000000, 4.000000000],
                             [5.000000000, 6.000000000, 7.000000000, 7.242640687, 5.806841831],
                             [5.801461031, 9.000000000, 12.000000000, 9.000000000, 6.074431804],
                             [5.895545844, 7.681980515, 9.000000000, 7.681980515, 5.935521488],
                             [4.988519641, 5.949655012, 6.170941546, 5.949655012, 4.997301087]])
    X = np.arange(25, dtype=float).reshape(5, 5)
    _X = X.copy()
    scarlet.operator.prox_soft_symmetry(_X, 0)
    assert_array_equal(_X, np.ones_like(X) * 12)
    _X = X.copy()
    scarlet.operator.prox_soft_symmetry(_X, 0, .5)
    assert_array_equal(_X[0], [6.0, 6.5, 7.0, 7.5, 8.0])
    x = np.zeros((21, 21))
    x[8:13, 8:13] = [[1, 2, 3, 2, 1], [2, 3, 4, 3, 1], [3, 4, 5, 1, 1], [2, 3, 1, 1, 1], [1, 1, 1, 1, 1]]
    _x = scarlet.operator.prox_kspace_symmetry(x, None, (0, 0))
    assert_almost_equal(_x, (x[::-1, ::-1] + x) / 2, decimal=6)
    X = np.arange(25, dtype=float).reshape(5, 5)
    scarlet.operator.prox_uncentered_symmetry(X, 0, (1, 1), "soft", None, None, 1)
    assert_array_equal(X[:3, :3], np.full((3, 3), 6.0))
    m = np.zeros((9, 9), np.float32)
    m[4, 4], m[2, 5], m[0, 0] = 1, 2, 10
    assert scarlet.measurement.max_pixel(m, (4, 4)) == (2, 5)


def test_update_surface(scarlet):
    """reference tests/test_update.py:120-211 style: component-level wrappers"""
    psf = scarlet.psf.generate_psf_image(scarlet.psf.gaussian, (15, 15), sigma=.9).image
    frame = scarlet.Frame((2, 5, 5), psfs=psf[None])
    morph = np.arange(25, dtype=np.float32).reshape(5, 5)
    c = scarlet.Component(frame, np.array([1., 2.]), morph.copy())
    scarlet.update.symmetric(c, (1, 1), algorithm="soft", strength=1)
    assert_array_equal(c.morph[:3, :3], np.full((3, 3), 6.0))
    assert_array_equal(c.morph[3:], morph[3:])
    c = scarlet.Component(frame, np.array([1., -2.]), morph.copy() - 3)
    scarlet.update.positive(c)
    assert c.sed[1] == 0 and c.morph.min() == 0
    scarlet.update.normalized(c)
    assert c.morph.max() == 1 and c.sed[0] == 21
    c = scarlet.Component(frame, np.array([1., 2.]), morph.copy())
    scarlet.update.monotonic(c, (2, 2), thresh=.25)
    assert_almost_equal(c.morph[1, 3:], [7.242640687, 5.806841831], decimal=5)


def test_two_observations_workflow(scarlet):
    """Blend(sources, [obs_a, obs_b]) with band-sliced observations (channels), as scarlet's multi-observation
    use: the fit through the Python surface follows the oracle's own two-observation loop."""
    rb = helpers.rb
    frame_o, many_o, srcs_o, blend_o = helpers.multi_obs_blend(7, kmin=3, kmax=4)
    C, Ny, Nx = frame_o.shape
    frame = scarlet.Frame(frame_o.shape, psfs=frame_o.psfs, channels=frame_o.channels)
    many = [scarlet.Observation(o.images, psfs=o.frame.psfs, weights=o.weights if isinstance(o.weights, np.ndarray) else None,
                                channels=o.frame.channels).match(frame) for o in many_o]
    assert [o._band_slice for o in many] == [o.band_slice for o in many_o]
    for a, b in zip(many, many_o):
        assert_allclose(a._diff_kernels.image, b.diff_kernels, rtol=0, atol=1e-7)
    srcs = []
    for so in srcs_o:
        s = scarlet.ExtendedSource(frame, so.pixel_center, None, None, _init=(so.sed.copy(), so.morph.copy()),
                                   _defer_update=True)
        s.pixel_center, s.shift = so.pixel_center, so.shift
        srcs.append(s)
    blend = scarlet.Blend(srcs, many)
    blend.fit(15, e_rel=0)
    for _ in range(15):
        blend_o.step(0.0)
    assert_allclose(blend.mse, blend_o.mse, rtol=2e-4)
    assert_array_equal([s.pixel_center for s in srcs], [s.pixel_center for s in srcs_o])
    gm = blend_o.get_model()
    assert_allclose(blend.get_model(), gm, rtol=0, atol=2e-4 * np.abs(gm).max())
    # per-observation post-fit products
    for o, oo in zip(many, many_o):
        assert_allclose(o.render(blend.get_model()), oo.render(gm), rtol=0, atol=3e-4 * np.abs(gm).max())


def test_large_frame_workflow(scarlet):
    """A 5-band 150x140 frame with 12 sources through the Python surface (config 4 class: beyond one convolution tile
    and beyond 8 components): ExtendedSource initialisation with the large-image operators, Blend.fit in scene mode,
    against the oracle's objects on the same data."""
    frame_o, obs_o, srcs_o, blend_o = helpers.synthetic_blend(21, Ny=150, Nx=140, kmin=10, kmax=12)
    frame = scarlet.Frame(frame_o.shape, psfs=frame_o.psfs)
    obs = scarlet.Observation(obs_o.images, psfs=obs_o.frame.psfs, weights=obs_o.weights).match(frame)
    bg = np.ones(frame.C, dtype=np.float32)
    srcs = [scarlet.ExtendedSource(frame, s.pixel_center_init, obs, bg) for s in srcs_o]
    blend = scarlet.Blend(srcs, obs)
    assert len(blend.components) == len(blend_o.components) > 8
    assert_array_equal([s.pixel_center for s in srcs], [s.pixel_center for s in srcs_o])
    assert_allclose(np.array([c.morph for c in blend.components]), np.array([c.morph for c in blend_o.components]),
                    rtol=0, atol=3e-6)
    blend.fit(8, e_rel=0)
    for _ in range(8):
        blend_o.step(0.0)
    assert_allclose(blend.mse, blend_o.mse, rtol=2e-4)
    gm = blend_o.get_model()
    assert_allclose(blend.get_model(), gm, rtol=0, atol=2e-4 * np.abs(gm).max())
    ren = obs.render(blend.get_model())                    # tiled device convolution
    assert_allclose(ren, obs_o.render(gm), rtol=0, atol=3e-4 * np.abs(gm).max())


def test_multiresolution_workflow(scarlet):
    """Frame(wcs) + Observation + LowResObservation + CombinedExtendedSource + Blend.fit, as in the reference's
    multi-resolution use, against the reference's own run of the same scene (tests/golden/make_golden_lowres.py)."""
    gold = np.load(os.path.join(GOLD, "ref_golden_lowres.npz"))
    helpers.sys_path_golden()
    from wcs_stub import make_pair
    for tag in ("A_", "B_"):
        hr_img, lr_img = gold[tag + "hr_img"], gold[tag + "lr_img"]
        whr, wlr = make_pair(hr_img.shape[1:], lr_img.shape[1:], scale=4.0, lr_offset=tuple(gold[tag + "offset"]))
        chan = ["h0", "h1", "h2", "l0", "l1", "l2", "l3", "l4"]
        frame = scarlet.Frame((8,) + hr_img.shape[1:], wcs=whr, psfs=gold[tag + "mpsf"][None], channels=chan)
        ohr = scarlet.Observation(hr_img, psfs=gold[tag + "hr_psfs"], wcs=whr, channels=chan[:3]).match(frame)
        olr = scarlet.LowResObservation(lr_img, wcs=wlr, psfs=gold[tag + "lr_psfs"], weights=gold[tag + "w_lr"],
                                        channels=chan[3:]).match(frame)
        assert list(olr._fft_shape) == list(gold[tag + "fft_shape"]) and olr.lr_shape == tuple(gold[tag + "lr_shape"])
        assert_array_equal(olr._coord_lr[0], gold[tag + "coord_lr0"])
        ref = gold[tag + "resconv_op"]
        assert_allclose(olr._resconv_op, ref, rtol=0, atol=2e-6 * np.abs(ref).max())
        ref = gold[tag + "render"]
        assert_allclose(olr._render(gold[tag + "model"]), ref, rtol=0, atol=1e-5 * np.abs(ref).max())
        assert_allclose(olr.get_loss(gold[tag + "model"]), gold[tag + "loss"], rtol=1e-5)
        assert olr.render(gold[tag + "model"]).shape == lr_img.shape
    # tag A: sources + fit
    tag = "A_"
    hr_img, lr_img = gold[tag + "hr_img"], gold[tag + "lr_img"]
    whr, wlr = make_pair(hr_img.shape[1:], lr_img.shape[1:], scale=4.0, lr_offset=tuple(gold[tag + "offset"]))
    frame = scarlet.Frame((8,) + hr_img.shape[1:], wcs=whr, psfs=gold[tag + "mpsf"][None], channels=chan)
    ohr = scarlet.Observation(hr_img, psfs=gold[tag + "hr_psfs"], wcs=whr, channels=chan[:3]).match(frame)
    olr = scarlet.LowResObservation(lr_img, wcs=wlr, psfs=gold[tag + "lr_psfs"], weights=gold[tag + "w_lr"],
                                    channels=chan[3:]).match(frame)
    HY, HX = hr_img.shape[1:]
    bg = [np.ones(3, np.float32), np.ones(5, np.float32)]
    srcs = [scarlet.CombinedExtendedSource(frame, (round(cy) - HY / 2, round(cx) - HX / 2), [ohr, olr], bg, obs_idx=0)
            for (cy, cx) in [(HY * .35, HX * .4), (HY * .6, HX * .62), (HY * .3, HX * .72)]]
    blend = scarlet.Blend(srcs, [ohr, olr])
    assert_array_equal([s.pixel_center for s in srcs], gold["A_init_centers"])
    assert_allclose(np.array([c.sed for c in blend.components]), gold["A_init_seds"], rtol=2e-6)
    assert_allclose(np.array([c.morph for c in blend.components]), gold["A_init_morphs"], rtol=0, atol=3e-6)
    ref = gold["A_mse"]
    blend.fit(len(ref), e_rel=1e-9)
    assert_allclose(blend.mse, ref, rtol=1e-4)
    assert_array_equal([s.pixel_center for s in srcs], gold["A_centers"])
    gm = gold["A_fit_model"]
    assert_allclose(blend.get_model(), gm, rtol=0, atol=1e-4 * np.abs(gm).max())


def test_unsupported_paths_are_loud(scarlet):
    frame, obs, srcs, blend = _scene(scarlet, "sim")
    # float32 and float64 frames have kernels (float64: tests/test_gpu_f64.py); anything else is refused
    f16 = scarlet.Frame(obs.images.shape, psfs=frame.psfs.image.astype(np.float16), dtype=np.float16)
    c = scarlet.Component(f16, np.ones(6), np.ones((40, 59)))
    with pytest.raises(NotImplementedError):
        scarlet.Blend([c], obs).fit(1)
