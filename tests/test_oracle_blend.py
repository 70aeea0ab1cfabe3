"""Oracle loop vs the reference's own loop (fixtures from tests/golden) and
vs torch.autograd for the analytic adjoint."""
import os
import numpy as np
import pytest
from numpy.testing import assert_array_equal, assert_allclose, assert_almost_equal

from oracle import ref_blend as rb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DATA = os.path.join(ROOT, "tests", "golden")


def _tb(golden):
    images = golden["tb_images"]
    frame = rb.Frame(images.shape, psfs=golden["tb_target_psf"][None], dtype=np.float64)
    obs = rb.Observation(images, psfs=golden["tb_psfs"]).match(frame)
    return frame, obs


COORDS = [(20, 10), (10, 30), (17, 42)]


def test_render_and_point_fit(golden):
    """reference tests/test_blend.py:56-94"""
    frame, obs = _tb(golden)
    assert_array_equal(obs.diff_kernels, golden["tb_diff_kernels"])
    blend = rb.Blend([rb.PointSource(frame, c, obs) for c in COORDS], obs)
    model = obs.render(blend.get_model())
    assert_array_equal(model, golden["tb_point_model"])
    assert_almost_equal(golden["tb_images"], model, decimal=5)
    blend.fit(10)
    assert blend.it == 2
    assert_almost_equal(blend.mse, [3.875628098330452e-15, 3.875598349723412e-15], decimal=10)
    assert blend.mse[0] > blend.mse[1]
    assert_allclose(blend.mse, golden["tb_point_mse"], rtol=1e-12)


def test_extended_fit(golden):
    """reference tests/test_blend.py:96-122"""
    frame, obs = _tb(golden)
    srcs = [rb.ExtendedSource(frame, c, obs, np.ones(6)) for c in COORDS]
    blend = rb.Blend(srcs, obs)
    assert_array_equal(np.array([c.sed for c in blend.components]), golden["tb_ext_init_seds"])
    assert_array_equal(np.array([c.morph for c in blend.components]), golden["tb_ext_init_morphs"])
    blend.fit(100)
    assert blend.it < 20
    mse = np.array(blend.mse)
    assert np.all(mse[:-1] - mse[1:] >= 0)
    assert len(mse) == len(golden["tb_ext_mse"])
    assert_allclose(mse, golden["tb_ext_mse"], rtol=1e-12)
    assert_allclose(np.array([c.morph for c in blend.components]), golden["tb_ext_morphs"], rtol=1e-10, atol=1e-14)
    assert_array_equal(np.array([s.pixel_center for s in srcs]), golden["tb_ext_centers"])


def _npz_blend(tag, golden, dtype=np.float32):
    fn = {"hsc": "hsc_cosmos_inputs.npz", "sim": "psf_unmatched_sim_inputs.npz"}[tag]
    d = np.load(os.path.join(DATA, fn))
    images, psfs = d["images"], d["psfs"]
    mpsf = rb.generate_psf_image(rb.gaussian, psfs[0].shape, sigma=.9)
    frame = rb.Frame(images.shape, psfs=mpsf[None], dtype=dtype)
    obs = rb.Observation(images, psfs=psfs).match(frame)
    bg = np.ones(len(images), dtype=images.dtype) * float(d["bg_rms"])
    srcs = []
    for (y, x), star in zip(d["coords"], d["is_star"]):
        if star:
            srcs.append(rb.PointSource(frame, (y, x), obs, fix_morph=True))
        else:
            srcs.append(rb.ExtendedSource(frame, (y, x), obs, bg))
    return frame, obs, srcs, rb.Blend(srcs, obs)


@pytest.mark.parametrize("tag", ["hsc", "sim"])
def test_npz_fit_vs_reference(golden, tag):
    """configs 1 and 2 (docs/quickstart.ipynb, docs/tutorials/point_source.ipynb)"""
    frame, obs, srcs, blend = _npz_blend(tag, golden)
    assert_array_equal(obs.diff_kernels, golden[tag + "_diff_kernels"])
    assert_array_equal(np.array([c.sed for c in blend.components]), golden[tag + "_init_seds"])
    assert_array_equal(np.array([c.morph for c in blend.components]), golden[tag + "_init_morphs"])
    assert_array_equal(np.array([s.pixel_center for s in srcs]), golden[tag + "_init_centers"])
    assert_array_equal(np.array([s.shift for s in srcs], dtype=np.float64), golden[tag + "_init_shifts"])
    blend.fit(200)
    ref_mse = golden[tag + "_mse"]
    assert blend.it == len(ref_mse)
    assert_allclose(np.array(blend.mse, dtype=np.float64), ref_mse, rtol=2e-6)
    assert_array_equal(np.array([s.pixel_center for s in srcs]), golden[tag + "_centers"])
    assert_allclose(blend.get_model(), golden[tag + "_model"], rtol=0,
                    atol=2e-5 * np.abs(golden[tag + "_model"]).max())


def _multi_blend(dtype=np.float32):
    """config 1 with the first rows as MultiComponentSource, like tests/golden/make_golden_multi.py"""
    d = np.load(os.path.join(DATA, "hsc_cosmos_inputs.npz"))
    g = np.load(os.path.join(DATA, "ref_golden_multi.npz"))
    images, psfs = d["images"], d["psfs"]
    mpsf = rb.generate_psf_image(rb.gaussian, psfs[0].shape, sigma=.9)
    frame = rb.Frame(images.shape, psfs=mpsf[None], dtype=dtype)
    obs = rb.Observation(images, psfs=psfs).match(frame)
    bg = np.ones(len(images), dtype=images.dtype) * float(d["bg_rms"])
    srcs = [(rb.MultiComponentSource if i < int(g["n_multi"]) else rb.ExtendedSource)(frame, (y, x), obs, bg)
            for i, (y, x) in enumerate(d["coords"])]
    return g, srcs, rb.Blend(srcs, obs)


def test_multicomponent_fit_vs_reference():
    """MultiComponentSource (source.py:537-641): init and the joint-centre update schedule inside the loop,
    against the reference itself (fixture ref_golden_multi.npz)."""
    g, srcs, blend = _multi_blend()
    assert [len(s.components) for s in srcs] == list(g["n_comp"])
    assert_allclose(np.array([c.sed for c in blend.components]), g["init_seds"], rtol=2e-6)
    assert_allclose(np.array([c.morph for c in blend.components]), g["init_morphs"], rtol=0, atol=2e-6)
    assert_array_equal(np.array([s.pixel_center for s in srcs]), g["init_centers"])
    blend.fit(int(g["max_iter"]))
    assert blend.it == len(g["mse"])
    assert_allclose(np.array(blend.mse, dtype=np.float64), g["mse"], rtol=5e-6)
    assert_array_equal(np.array([s.pixel_center for s in srcs]), g["centers"])
    assert_allclose(blend.get_model(), g["model"], rtol=0, atol=2e-5 * np.abs(g["model"]).max())


def test_adjoint_vs_torch_autograd(golden):
    """The analytic gradient that replaces autograd.grad(Blend._loss) (blend.py:44-45)."""
    import torch
    frame, obs = _tb(golden)
    rng = np.random.default_rng(0)
    K, (C, Ny, Nx) = 3, frame.shape
    comps = [rb.Component(frame, rng.random(C) + .1, rng.random((Ny, Nx))) for _ in range(K)]
    obs.weights = rng.random(frame.shape)
    blend = rb.Blend(comps, obs)
    sg, mg = blend.backward()

    from oracle import ref_fft
    fshape = ref_fft.fft_shape(frame.shape, obs.diff_kernels.shape, 3, (1, 2))
    D = torch.from_numpy(ref_fft.forward(obs.diff_kernels, fshape, (1, 2)))
    A = torch.tensor(np.array([c.sed for c in comps]), requires_grad=True)
    S = torch.tensor(np.array([c.morph for c in comps]), requires_grad=True)
    M = torch.einsum("kc,kyx->cyx", A, S)
    ly, lx = [(f - n + 1) // 2 for f, n in zip(fshape, (Ny, Nx))]
    P = torch.nn.functional.pad(M, (lx, fshape[1] - Nx - lx, ly, fshape[0] - Ny - ly))
    spec = torch.fft.rfftn(torch.fft.ifftshift(P, dim=(1, 2)), dim=(1, 2)) * D
    R = torch.fft.fftshift(torch.fft.irfftn(spec, s=fshape, dim=(1, 2)), dim=(1, 2))[:, ly:ly + Ny, lx:lx + Nx]
    loss = 0.5 * torch.sum((torch.from_numpy(obs.weights) * (R - torch.from_numpy(obs.images))) ** 2)
    loss.backward()
    assert_allclose(loss.item(), blend.mse[0], rtol=1e-12)
    assert_allclose(np.array(sg), A.grad.numpy(), rtol=1e-9)
    assert_allclose(np.array(mg), S.grad.numpy(), rtol=1e-8, atol=1e-12 * np.abs(S.grad.numpy()).max())
