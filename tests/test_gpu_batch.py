"""GPU parity of BlendBatch (array-level batch entry, config 3) against the oracle, plus size-independent
properties at a larger batch."""
import numpy as np
import pytest
from numpy.testing import assert_array_equal, assert_allclose

import helpers

pytestmark = pytest.mark.gpu
REL = 1e-5


@pytest.fixture(scope="module")
def batch6():
    from scarlet_b200 import synth
    from scarlet_b200.batch import BlendBatch
    d = synth.make_blends(range(6), device="cpu")
    bb = BlendBatch(d["images"], d["psfs"], d["model_psf"], d["centers"], d["bg_rms"], weights=d["weights"])
    yield d, bb
    bb.close()


def test_init_matches_oracle_sources(batch6):
    """Source initialisation + the construction-time constraint pass equal the oracle's ExtendedSource objects."""
    d, bb = batch6
    off = bb.host["comp_off"]
    for i in range(6):
        _, obs, srcs, _ = helpers.synthetic_blend(i)
        assert_array_equal(obs.images, d["images"][i])
        assert off[i + 1] - off[i] == len(srcs)
        for j, s in enumerate(srcs):
            k = off[i] + j
            assert_array_equal(bb.host["center"][k], s.pixel_center)
            assert_allclose(bb.host["shift"][k], s.shift, rtol=0, atol=5e-6)
            assert_allclose(bb.host["sed"][k], s.sed, rtol=2e-6)
            assert_allclose(bb.host["morph"][k], s.morph, rtol=0, atol=REL)


def _oracle_sources(d, i, kinds, dtype=np.float32):
    """Oracle sources of blend i of a synth.make_blends dict, one per catalogue row, of the given kinds."""
    from oracle import ref_blend as rb
    frame = rb.Frame(d["images"][i].shape, psfs=d["model_psf"], dtype=dtype)
    obs = rb.Observation(d["images"][i], psfs=d["psfs"], weights=d["weights"][i]).match(frame)
    srcs = []
    for (y, x), kd in zip(d["centers"][i], kinds):
        if kd == 0:
            srcs.append(rb.ExtendedSource(frame, (int(y), int(x)), obs, d["bg_rms"]))
        elif kd == 1:
            srcs.append(rb.MultiComponentSource(frame, (int(y), int(x)), obs, d["bg_rms"]))
        else:
            srcs.append(rb.PointSource(frame, (int(y), int(x)), obs))
    return frame, obs, srcs


@pytest.mark.parametrize("shape", [(64, 64), (150, 140)])
def test_init_of_every_source_kind_matches_oracle(shape):
    """ExtendedSource, MultiComponentSource (flux layers + least-squares SEDs, source.py:74-98,242-291)
    and PointSource (source.py:355-375) initialised on the device, then the construction-time update(); against the
    oracle's objects.  (150, 140): the same in scene mode (working copies in the HBM scratch), then three iterations of
    the mixed batch in lock step."""
    from scarlet_b200 import synth
    from scarlet_b200.batch import BlendBatch
    from scarlet_b200 import _native as nat
    from oracle import ref_blend as rb
    n = 4 if shape == (64, 64) else 1
    d = synth.make_blends(range(20, 20 + n), Ny=shape[0], Nx=shape[1], kmin=4, kmax=6, device="cpu")
    kinds = [np.arange(len(c)) % 3 for c in d["centers"]]
    bb = BlendBatch(d["images"], d["psfs"], d["model_psf"], d["centers"], d["bg_rms"], weights=d["weights"], kinds=kinds)
    assert bb.n_failed == 0
    off = bb.host["comp_off"]
    blends = []
    for i in range(n):
        frame, obs, srcs = _oracle_sources(d, i, kinds[i])
        k = off[i]
        for s in srcs:
            for c in s.components:
                assert_array_equal(bb.host["center"][k], s.pixel_center)
                assert_allclose(bb.host["shift"][k], s.shift, rtol=0, atol=5e-6)
                assert_allclose(bb.host["sed"][k], c.sed, rtol=2e-4, atol=2e-5 * np.abs(c.sed).max())
                assert_allclose(bb.host["morph"][k], c.morph, rtol=0, atol=REL)
                k += 1
        assert k == off[i + 1]
        blends.append(rb.Blend(srcs, obs))
    # the loop on the mixed batch, from the oracle's own initial state, three iterations
    import test_gpu_fit
    for bl in blends[:2]:
        test_gpu_fit._lockstep(nat, bl, 3)
    bb.close()


def test_fit_matches_oracle(batch6):
    """Free-running fit(30, e_rel=0) of the batch vs the oracle blend by blend."""
    d, bb = batch6
    bb.snapshot()
    bb.fit_device(30, 0.0)
    out = bb.download()
    off = bb.host["comp_off"]
    for i in range(6):
        bl = helpers.synthetic_blend(i)[3]
        bl.fit(30, 0.0)
        assert out["n_iter"][i] == 30
        assert_allclose(out["logL"][i], np.array(bl.mse, dtype=np.float64), rtol=20 * REL)
        assert_array_equal(out["center"][off[i]:off[i + 1]], [s.pixel_center for s in bl.sources])
        ref = bl.get_model().astype(np.float64)
        got = np.einsum("kc,kyx->cyx", out["sed"][off[i]:off[i + 1]].astype(np.float64),
                        out["morph"][off[i]:off[i + 1]].astype(np.float64))
        assert_allclose(got, ref, rtol=0, atol=2e-4 * np.abs(ref).max())    # 30 free-running fp32 iterations; single iterations: test_gpu_parity_config3.py
    # restore brings the initial state back bit for bit; the repeated fit is bit-identical
    bb.restore()
    again0 = bb.download()
    assert_array_equal(again0["morph"], bb.host["morph"])
    assert_array_equal(again0["sed"], bb.host["sed"])
    bb.fit_device(30, 0.0)
    again = bb.download()
    assert_array_equal(again["logL"], out["logL"])
    assert_array_equal(again["morph"], out["morph"])
    bb.restore()


def test_fit_host_roundtrip_and_converged_stop(batch6):
    """BlendBatch.fit (upload + loop + download) with the reference defaults: every blend stops on its own
    iteration, the loss history is non-increasing at the end and the state after a converged stop is frozen."""
    d, bb = batch6
    info = bb.fit(200, 1e-2)
    assert info["launches"] > 0
    out = bb.last
    assert (out["active"] == 0).all() and (out["conv"] == 0).all()
    for i in range(bb.B):
        bl = helpers.synthetic_blend(i)[3]
        bl.fit(200, 1e-2)
        assert abs(len(bb.mse[i]) - bl.it) <= 1          # a 1-ulp difference may move the stopping iteration by one
        n = min(len(bb.mse[i]), bl.it)
        assert_allclose(bb.mse[i][:n], np.array(bl.mse[:n], dtype=np.float64), rtol=20 * REL)


def test_properties_large_batch():
    """512 blends (BASELINE shape): per-blend results do not depend on the batch they run in; morphologies obey the
    constraints (non-negative, peak-normalised, monotonic from the peak outwards)."""
    from scarlet_b200 import synth
    from scarlet_b200.batch import BlendBatch
    d = synth.make_blends(range(100, 612), device="cpu")
    bb = BlendBatch(d["images"], d["psfs"], d["model_psf"], d["centers"], d["bg_rms"], weights=d["weights"])
    bb.fit_device(200, 1e-2)
    out = bb.download()
    off = bb.host["comp_off"]
    assert (out["active"] == 0).mean() > 0.9
    assert (out["morph"] >= 0).all()
    assert_allclose(out["morph"].reshape(bb.K, -1).max(axis=1), 1.0, rtol=1e-6)
    peak = out["morph"].reshape(bb.K, -1).argmax(axis=1)
    # the peak sits within the 5x5 window of the tracked centre
    cy, cx = out["center"][:, 0], out["center"][:, 1]
    assert (np.abs(peak // 64 - cy) <= 2).all() and (np.abs(peak % 64 - cx) <= 2).all()
    # loss never increases over the last iterations of any blend
    for b in range(0, bb.B, 37):
        n = out["n_iter"][b]
        l = out["logL"][b, :n]
        assert l[-1] <= l[0]
    # sub-batch independence: blends 40..47 alone give bit-identical results
    sel = list(range(140, 148))
    d2 = synth.make_blends(sel, device="cpu")
    b2 = BlendBatch(d2["images"], d2["psfs"], d2["model_psf"], d2["centers"], d2["bg_rms"], weights=d2["weights"])
    b2.fit_device(200, 1e-2)
    o2 = b2.download()
    kept = list(bb.kept)
    assert b2.B == len(sel)
    for j, g in enumerate(sel):
        i = kept.index(g - 100)
        assert out["n_iter"][i] == o2["n_iter"][j]
        assert_array_equal(out["logL"][i, :out["n_iter"][i]], o2["logL"][j, :o2["n_iter"][j]])
        assert_array_equal(out["morph"][off[i]:off[i + 1]], o2["morph"][b2.host["comp_off"][j]:b2.host["comp_off"][j + 1]])
    bb.close()
    b2.close()


def test_chunked_streams_equal_single_batch():
    """A batch split into independently driven chunks (one CUDA stream + host thread each) gives bit-identical
    results to the single-handle batch: blends never interact, whatever the chunk boundaries."""
    from scarlet_b200 import synth
    from scarlet_b200.batch import BlendBatch
    d = synth.make_blends(range(23), device="cpu")
    args = (d["images"], d["psfs"], d["model_psf"], d["centers"], d["bg_rms"])
    one = BlendBatch(*args, weights=d["weights"])
    many = BlendBatch(*args, weights=d["weights"], streams=4)
    assert len(many._dev.parts) == 4
    for k in ("sed", "morph", "center", "shift", "conv"):
        assert_array_equal(one.host[k], many.host[k])
    i1 = one.fit(25, 1e-2)
    i2 = many.fit(25, 1e-2)
    assert i2["launches"] >= i1["launches"]
    for k in ("sed", "morph", "center", "shift", "conv", "n_iter", "active", "lipschitz"):
        assert_array_equal(one.last[k], many.last[k])
    n = one.last["n_iter"]
    for b in range(one.B):
        assert_array_equal(one.last["logL"][b, :n[b]], many.last["logL"][b, :n[b]])
        assert one.mse[b] == many.mse[b]
    # a second call continues from the first on both
    one.fit(5, 0.0)
    many.fit(5, 0.0, )
    assert_array_equal(one.host["morph"], many.host["morph"])
    # snapshot / restore / serial mode
    many.snapshot()
    many.fit_device(3, 0.0)
    many.restore()
    a = many.fit_device(3, 0.0, serial=True)
    assert a["steps"] == 3
    one.close()
    many.close()


def test_failed_sources_are_dropped_like_source_init_error():
    """A catalogue position with no pixel above the noise cut raises SourceInitError in the reference
    (source.py:171-173); the batch drops such sources (and blends left empty) and reports them, and the surviving
    sources are initialised exactly as in a batch that never contained the bad ones."""
    from scarlet_b200 import synth
    from scarlet_b200.batch import BlendBatch
    from oracle import ref_blend as rb
    d = synth.make_blends(range(4), device="cpu")
    images = d["images"].copy()
    centers = [np.asarray(c).copy() for c in d["centers"]]
    images[3] = 0.0                                            # blend 3: S/N << 1 at every catalogue position
    images[3][:, centers[3][:, 0], centers[3][:, 1]] = 1e-3
    # blend 1 gets one extra catalogue row on a faint corner pixel: monotonicity from there leaves nothing above the cut
    images[1, :, 2, 2] = 1e-3
    centers[1] = np.concatenate([centers[1], [[2, 2]]])
    args = (d["psfs"], d["model_psf"])
    bb = BlendBatch(images, *args, centers, d["bg_rms"], weights=d["weights"])
    assert bb.n_failed == 1 + len(centers[3])
    assert list(bb.kept) == [0, 1, 2]
    # oracle: the bad row raises, the others build
    frame = rb.Frame(images[1].shape, psfs=np.asarray(d["model_psf"]), dtype=np.float32)
    obs = rb.Observation(images[1], psfs=np.asarray(d["psfs"])).match(frame)
    with pytest.raises(rb.SourceInitError):
        rb.ExtendedSource(frame, (2, 2), obs, np.asarray(d["bg_rms"], dtype=np.float32))
    ref = BlendBatch(images[:3], *args, [centers[0], centers[1][:-1], centers[2]], d["bg_rms"], weights=d["weights"][:3])
    assert ref.n_failed == 0
    for k in ("comp_off", "sed", "morph", "center", "shift"):
        assert_array_equal(bb.host[k], ref.host[k])
    bb.close()
    ref.close()


def test_products_of_a_batch():
    """BlendBatch.products / render_components on the fitted batch: fluxes, footprint boxes and flags
    against numpy on the downloaded state, per-component renders summing to the oracle's rendered blend model; the
    chunked batch returns the same arrays."""
    from scarlet_b200 import synth
    from scarlet_b200.batch import BlendBatch
    d = synth.make_blends(range(6), device="cpu")
    bb = BlendBatch(d["images"], d["psfs"], d["model_psf"], d["centers"], d["bg_rms"], weights=d["weights"])
    bb.fit(15, 0.0)
    flux, bounds, flags = bb.products()
    ren = bb.render_components()
    off = bb.host["comp_off"]
    morph = bb.host["morph"].reshape(len(flux), *bb.shape[1:])
    assert_allclose(flux, morph.sum(axis=(1, 2))[:, None] * bb.host["sed"], rtol=2e-6)
    for k in range(len(flux)):
        nz = np.where(morph[k] > 0)
        assert_array_equal(bounds[k], [nz[0].min(), nz[0].max(), nz[1].min(), nz[1].max()])
    assert np.all((flags & 3) == (bb.host["conv"] & 3)) and not np.any(flags & 8)
    for i in range(6):
        obs = helpers.synthetic_blend(i)[1]
        want = obs.render(bb.get_model(i).astype(np.float64))
        assert_allclose(ren[off[i]:off[i + 1]].sum(axis=0), want, rtol=0, atol=2e-5 * np.abs(want).max())
    many = BlendBatch(d["images"], d["psfs"], d["model_psf"], d["centers"], d["bg_rms"], weights=d["weights"], streams=3)
    many.fit(15, 0.0)
    f2, b2, g2 = many.products()
    assert_array_equal(f2, flux)
    assert_array_equal(b2, bounds)
    assert_array_equal(g2, flags)
    assert_array_equal(many.render_components(2, 9), ren[2:11])
    many.close()
    bb.close()


def test_per_blend_psfs_matched_inside_the_batch():
    """Every blend with its own observed PSFs (BASELINE config 2 at scale): Observation.match runs on the device inside
    the batch (scl_batch_desc.psf_obs) and the spectra stay in HBM.  (a) the shared PSFs tiled per blend give the
    shared-PSF batch bit for bit, chunked or not; (b) four blends with four different PSF sets follow their oracle
    blends (initialisation and 12 free-running iterations)."""
    from scarlet_b200 import synth
    from scarlet_b200.batch import BlendBatch
    from oracle import ref_blend as rb
    d = synth.make_blends(range(5), device="cpu")
    B, C = d["images"].shape[:2]
    args = (d["model_psf"], d["centers"], d["bg_rms"])
    one = BlendBatch(d["images"], d["psfs"], *args, weights=d["weights"])
    tiled = np.broadcast_to(np.asarray(d["psfs"], np.float32), (B,) + np.shape(d["psfs"])).copy()
    per = BlendBatch(d["images"], tiled, *args, weights=d["weights"])
    chunked = BlendBatch(d["images"], tiled, *args, weights=d["weights"], streams=2)
    for bb in (one, per, chunked):
        bb.fit(12, 0.0)
    for k in ("sed", "morph", "center", "shift", "logL"):
        assert_array_equal(per.last[k], one.last[k])
        assert_array_equal(chunked.last[k], one.last[k])
    for bb in (one, per, chunked):
        bb.close()

    ps = np.shape(d["psfs"])[-1]
    sets = np.array([[rb.generate_psf_image(rb.gaussian, (ps, ps), sigma=1.15 + .1 * i + .15 * b) for b in range(C)]
                     for i in range(4)], dtype=np.float32)
    bb = BlendBatch(d["images"][:4], sets, *(d["model_psf"], d["centers"][:4], d["bg_rms"]), weights=d["weights"][:4])
    init = {k: np.array(v) for k, v in bb.host.items() if k in ("sed", "morph", "center")}
    bb.fit(12, 0.0)
    off = bb.host["comp_off"]
    mpsf = np.asarray(d["model_psf"], np.float32)
    for i in range(4):
        frame = rb.Frame(d["images"][i].shape, psfs=mpsf, dtype=np.float32)
        obs = rb.Observation(d["images"][i], psfs=sets[i], weights=d["weights"][i]).match(frame)
        srcs = [rb.ExtendedSource(frame, tuple(int(v) for v in c), obs, np.asarray(d["bg_rms"], np.float32))
                for c in np.asarray(d["centers"][i]).reshape(-1, 2)]
        assert off[i + 1] - off[i] == len(srcs)
        for j, s in enumerate(srcs):
            assert_array_equal(init["center"][off[i] + j], s.pixel_center)
            assert_allclose(init["sed"][off[i] + j], s.sed, rtol=2e-6)
            assert_allclose(init["morph"][off[i] + j].reshape(s.morph.shape), s.morph, rtol=0, atol=REL)
        bl = rb.Blend(srcs, obs)
        bl.fit(12, 0.0)
        assert_allclose(bb.last["logL"][i], np.array(bl.mse, dtype=np.float64), rtol=20 * REL)
    bb.close()


def test_chunked_batch_on_second_device():
    """The chunk worker threads adopt the device of the thread that built the batch (the CUDA device is a per-thread
    setting; rank r of a multi-GPU job runs on cuda:r).  Needs two GPUs; runs in a fresh process because the
    library keeps its tables on the first device it is used on (one process per GPU)."""
    import subprocess
    import sys
    from scarlet_b200 import _native
    if _native.lib().scl_device_count() < 2:
        pytest.skip("needs two GPUs")
    code = (
        "import numpy as np\n"
        "from scarlet_b200 import _native, synth\n"
        "from scarlet_b200.batch import BlendBatch\n"
        "_native.device_set(1)\n"
        "d = synth.make_blends(range(9), device='cpu')\n"
        "a = (d['images'], d['psfs'], d['model_psf'], d['centers'], d['bg_rms'])\n"
        "one = BlendBatch(*a, weights=d['weights']); many = BlendBatch(*a, weights=d['weights'], streams=3)\n"
        "one.fit(10, 0.0); many.fit(10, 0.0)\n"
        "assert np.array_equal(one.last['morph'], many.last['morph']) and np.array_equal(one.last['logL'], many.last['logL'])\n"
        "print('ok')\n")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600,
                       cwd=helpers.ROOT if hasattr(helpers, "ROOT") else None)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stderr[-2000:]
