"""Blend.fit on float64 frames (scl_fit_f64, csrc/k_f64.cu): lock-step against the float64 oracle, and the bodies of
the reference's own loop tests (/root/reference/tests/test_blend.py:7-122, which build Frame(dtype=np.float64))
through `import scarlet`."""
import numpy as np
import pytest
from numpy.testing import assert_array_equal, assert_allclose, assert_almost_equal

import helpers

pytestmark = pytest.mark.gpu

REL = 1e-10     # float64 against the float64 oracle (whose convolution is an FFT product: ~1e-13 of rounding)


@pytest.fixture(scope="module")
def nat():
    from scarlet_b200 import _native
    _native.require_device()
    return _native


@pytest.fixture(scope="module")
def scarlet():
    import scarlet
    scarlet._native.require_device()
    return scarlet


def _args64(blend):
    """Arguments of _native.fit_f64 from an oracle Blend (float64 frame)."""
    p = helpers.pack([blend])
    C, Ny, Nx = blend.frame.shape
    kernels, ker_on = [], []
    ks = [o.diff_kernels.shape[1:] for o in blend.observations if o.diff_kernels is not None]
    for o in blend.observations:
        Co = len(o.images)
        if o.diff_kernels is None:
            ker_on += [0] * Co
            kernels += [np.zeros(ks[0])] * Co if ks else []
        else:
            ker_on += [1] * Co
            kernels += list(o.diff_kernels)
    plane_band = p.get("plane_band", np.arange(C))
    images = np.concatenate([o.images for o in blend.observations]).astype(np.float64)
    any_w = any(isinstance(o.weights, np.ndarray) for o in blend.observations)
    weights = np.concatenate([np.broadcast_to(o.weights, o.images.shape) for o in blend.observations]) if any_w else None
    return dict(C=C, Ny=Ny, Nx=Nx, sed=np.array([c.sed for c in blend.components]),
                morph=np.array([c.morph for c in blend.components]), cflags=p["cflags"], center=p["center"],
                shift=p["shift"], conv=p["conv"], images=images, weights=weights, plane_band=plane_band,
                n_obs=len(blend.observations), kernels=np.array(kernels) if ks else None, ker_on=ker_on,
                cw=helpers.rb._centroid_weight(blend.frame), it0=blend.it, prev_loss=blend.mse[-1] if blend.mse else 0.,
                src_first=p["src_first"])


def _lockstep64(nat, blend, n_steps, e_rel=1e-2, approximate_L=False):
    for _ in range(n_steps):
        out = nat.fit_f64(max_iter=1, e_rel=e_rel, approximate_L=approximate_L, **_args64(blend))
        conv = blend.step(e_rel, approximate_L=approximate_L)
        assert out["n_iter"] == 1
        assert_allclose(out["logL"][0], blend.mse[-1], rtol=REL)
        assert_allclose(out["lipschitz"], [blend.L_sed, blend.L_morph], rtol=REL)
        con = np.array([isinstance(s, (helpers.rb.PointSource, helpers.rb.MultiComponentSource))
                        for s in blend.sources for _ in s.components])
        ref_c = np.array([getattr(s, "pixel_center", (0, 0)) for s in blend.sources for _ in s.components])
        assert_array_equal(out["center"][con], ref_c[con])
        ref_s = np.array([getattr(s, "shift", (0, 0)) for s in blend.sources for _ in s.components], dtype=np.float64)
        assert_allclose(out["shift"][con], ref_s[con], rtol=0, atol=1e-9)
        sed = np.array([c.sed for c in blend.components])
        morph = np.array([c.morph for c in blend.components])
        assert_allclose(out["sed"], sed, rtol=0, atol=REL * np.abs(sed).max())
        assert_allclose(out["morph"], morph, rtol=0, atol=REL)
        assert_array_equal(out["conv"], [c.flags for c in blend.components])
        assert out["converged"] == bool(conv)
        if conv:
            break


@pytest.mark.parametrize("approx", [False, True])
@pytest.mark.parametrize("tag", ["hsc", "sim"])
def test_lockstep_f64_npz(nat, tag, approx):
    """configs 1 and 2 with float64 frames, every iteration vs the float64 oracle."""
    _, _, _, blend = helpers.npz_blend(tag, dtype=np.float64)
    assert blend.components[0].morph.dtype == np.float64
    _lockstep64(nat, blend, 10, approximate_L=approx)


def test_lockstep_f64_multicomponent(nat):
    _, _, _, _, blend = helpers.multi_blend(dtype=np.float64)
    _lockstep64(nat, blend, 8)


def test_lockstep_f64_two_observations(nat):
    _, _, _, blend = helpers.multi_obs_blend(3, kmin=3, kmax=5, dtype=np.float64)
    _lockstep64(nat, blend, 5)


def test_free_running_f64_vs_f32_reference_run(nat, golden):
    """The whole float64 fit of config 1 against the reference's own float32 run (fixture): same iteration count,
    logL within float32 rounding."""
    _, _, _, blend = helpers.npz_blend("hsc", dtype=np.float64)
    out = nat.fit_f64(max_iter=200, e_rel=1e-2, **_args64(blend))
    ref = golden["hsc_mse"]
    assert abs(out["n_iter"] - len(ref)) <= 1 and out["converged"]
    n = min(out["n_iter"], len(ref))
    assert_allclose(out["logL"][:n], ref[:n], rtol=2e-4)


# ------------------------------------------------------------------ the reference's tests/test_blend.py, restated
def init_data(scarlet, shape, coords, amplitudes=None, dtype=np.float32):
    """tests/test_blend.py:7-52"""
    import scipy.signal
    B, Ny, Nx = shape
    K = len(coords)
    if amplitudes is None:
        amplitudes = np.ones((K,))
    _seds = [np.arange(B, dtype=dtype), np.arange(B, dtype=dtype)[::-1], np.ones((B,), dtype=dtype)]
    seds = np.array([_seds[n % 3] * amplitudes[n] for n in range(K)])
    morphs = np.zeros((K, Ny, Nx))
    for k, coord in enumerate(coords):
        morphs[k, coord[0], coord[1]] = 1
    images = seds.T.dot(morphs.reshape(K, -1)).reshape(shape)
    psf_shape = (41, 41)
    target_psf = scarlet.psf.generate_psf_image(scarlet.psf.gaussian, psf_shape, sigma=.9).image
    target_psf /= target_psf.sum()
    psfs = np.array([scarlet.psf.generate_psf_image(scarlet.psf.gaussian, psf_shape, sigma=1 + .2 * b).image
                     for b in range(B)], dtype=dtype)
    psfs /= psfs.max(axis=(1, 2))[:, None, None]
    images = np.array([scipy.signal.convolve(img, psf, method="direct", mode="same")
                       for img, psf in zip(images, psfs)], dtype=dtype)
    morphs = np.array([scipy.signal.convolve(m, target_psf, method="direct", mode="same") for m in morphs], dtype=dtype)
    morphs /= morphs.max()
    psfs /= psfs.sum(axis=(1, 2))[:, None, None]
    return target_psf, psfs, images, range(len(images)), seds, morphs


SHAPE, COORDS, AMPS = (6, 31, 55), [(20, 10), (10, 30), (17, 42)], [3, 2, 1]


def test_reference_model_render(scarlet):
    """tests/test_blend.py:56-73"""
    target_psf, psfs, images, channels, seds, morphs = init_data(scarlet, SHAPE, COORDS, AMPS, dtype=np.float64)
    frame = scarlet.Frame(images.shape, psfs=target_psf[None], dtype=np.float64)
    observation = scarlet.Observation(images, psfs=psfs).match(frame)
    sources = [scarlet.PointSource(frame, coord, observation) for coord in COORDS]
    blend = scarlet.Blend(sources, observation)
    model = observation.render(blend.get_model())
    assert model.dtype == np.float64
    assert_almost_equal(images, model, decimal=5)
    for s0, s in zip(sources, blend.sources):
        assert_array_equal(s.get_model(), s0.get_model())


def test_reference_fit_point_source(scarlet):
    """tests/test_blend.py:75-94"""
    target_psf, psfs, images, channels, seds, morphs = init_data(scarlet, SHAPE, COORDS, AMPS, dtype=np.float64)
    frame = scarlet.Frame(images.shape, psfs=target_psf[None], dtype=np.float64)
    observation = scarlet.Observation(images, psfs=psfs).match(frame)
    sources = [scarlet.PointSource(frame, coord, observation) for coord in COORDS]
    blend = scarlet.Blend(sources, observation)
    blend.fit(10)
    assert blend.it == 2
    assert_almost_equal(blend.mse, [3.875628098330452e-15, 3.875598349723412e-15], decimal=10)
    assert blend.mse[0] > blend.mse[1]
    assert blend.components[0].morph.dtype == np.float64


def test_reference_fit_extended_source(scarlet):
    """tests/test_blend.py:96-122"""
    target_psf, psfs, images, channels, seds, morphs = init_data(scarlet, SHAPE, COORDS, AMPS, dtype=np.float64)
    B = SHAPE[0]
    frame = scarlet.Frame(images.shape, psfs=target_psf[None], dtype=np.float64)
    observation = scarlet.Observation(images, psfs=psfs).match(frame)
    bg_rms = np.ones((B,))
    sources = [scarlet.ExtendedSource(frame, coord, observation, bg_rms) for coord in COORDS]
    blend = scarlet.Blend(sources, observation)
    psf_scale = observation.frame.psfs.max(axis=(1, 2)) / frame.psfs[0].max()
    scaled_seds = np.array([c.sed * psf_scale for c in blend.components])
    assert_almost_equal(scaled_seds, seds)
    blend.fit(100)
    assert blend.it < 20
    mse = np.array(blend.mse[:-1])
    _mse = np.array(blend.mse[1:])
    assert np.all(mse - _mse >= 0)
