"""GPU parity of the fused Blend.fit loop (through the C ABI) against the oracle."""
import copy
import numpy as np
import pytest
from numpy.testing import assert_array_equal, assert_allclose

import helpers

pytestmark = pytest.mark.gpu

REL = 1e-5      # north_star tolerance on per-iteration logL and on the model (fp32)


@pytest.fixture(scope="module")
def nat():
    from scarlet_b200 import _native
    _native.require_device()
    return _native


def _model(out, k0, k1):
    return np.einsum("kc,kyx->cyx", out["sed"][k0:k1].astype(np.float64), out["morph"][k0:k1].astype(np.float64))


def _lockstep(nat, blend, n_steps, e_rel=1e-2, approximate_L=False, **batch_kw):
    """Each iteration: start the GPU from the oracle's state, run one iteration on both,
    compare everything the iteration produces."""
    for step in range(n_steps):
        p = helpers.pack([blend])
        p.update(batch_kw)
        bt = nat.Batch(**p)
        bt.fit(1, e_rel, approximate_L=approximate_L)
        out = bt.download()
        bt.close()
        conv = blend.step(e_rel, approximate_L=approximate_L)
        assert_allclose(out["logL"][0, 0], blend.mse[-1], rtol=REL)
        assert_allclose(out["lipschitz"][0], [blend.L_sed, blend.L_morph], rtol=2e-6)
        ref_c = np.array([getattr(s, "pixel_center", (0, 0)) for s in blend.sources for _ in s.components])
        con = np.array([isinstance(s, (helpers.rb.PointSource, helpers.rb.MultiComponentSource))
                        for s in blend.sources for _ in s.components])
        assert_array_equal(out["center"][con], ref_c[con])                      # index work: exact
        ref_s = np.array([getattr(s, "shift", (0, 0)) for s in blend.sources for _ in s.components], dtype=np.float64)
        assert_allclose(out["shift"][con], ref_s[con], rtol=0, atol=5e-6)
        sed = np.array([c.sed for c in blend.components])
        morph = np.array([c.morph for c in blend.components])
        assert_allclose(out["sed"], sed, rtol=0, atol=REL * np.abs(sed).max())
        assert_allclose(out["morph"], morph, rtol=0, atol=REL)
        assert_array_equal(out["conv"], [c.flags for c in blend.components])
        assert out["active"][0] == (0 if conv else 1)
        if conv:
            break


@pytest.mark.parametrize("tag", ["hsc", "sim"])
def test_lockstep_npz(nat, tag):
    """configs 1 and 2, every iteration until the oracle converges"""
    _, _, _, blend = helpers.npz_blend(tag)
    _lockstep(nat, blend, 60)


def test_lockstep_multicomponent(nat):
    """MultiComponentSource.update inside the loop (source.py:605-641): the joint centre of a source is measured on
    the flux-weighted sum of its stepped components (scl_k_source_center), every iteration vs the oracle."""
    _, _, _, _, blend = helpers.multi_blend()
    _lockstep(nat, blend, 40)


@pytest.mark.parametrize("split", [((0, 3), (1, 5)), ((0, 4), (1, 5)), ((2, 5), (0, 2), (1, 4))])
def test_lockstep_multiple_observations(nat, split):
    """Several band-sliced observations of one blend (Blend._loss sums their losses and gradients, blend.py:135-139;
    both Lipschitz constants scale with len(observations), blend.py:219-220): stacked image planes on the device."""
    if sum(hi - lo for lo, hi in split) > 8:
        pytest.skip("more than 8 planes")
    _, obs, _, blend = helpers.multi_obs_blend(3, split=split, kmin=3, kmax=5)
    assert len(blend.observations) == len(split)
    _lockstep(nat, blend, 12)


def test_multiple_observations_approximate_L(nat):
    """approximate_L with two observations: free-running logL against the oracle"""
    _, _, _, blend = helpers.multi_obs_blend(5, kmin=2, kmax=4)
    p = helpers.pack([blend])
    bt = nat.Batch(**p)
    bt.fit(10, 0.0, approximate_L=True)
    out = bt.download()
    bt.close()
    for _ in range(10):
        blend.step(0.0, approximate_L=True)
    assert_allclose(out["logL"][0, :10], blend.mse, rtol=2e-4)


def test_lockstep_rotated_low_resolution_observation(nat):
    """The same with the low-resolution frame rotated by 10 degrees against the model frame (observation.py:443-454,
    483-503; fixture R_ of a run of the reference): every iteration against the oracle's restated rotated `_render`."""
    import os
    gold = np.load(os.path.join(helpers.GOLD, "ref_golden_lowres.npz"))
    frame, ohr, olr = helpers.lowres_scene(gold, "R_")
    assert olr.isrot
    srcs, blend = helpers.lowres_blend(gold, frame, ohr, olr)
    _lockstep(nat, blend, 6, e_rel=1e-9)


def test_lockstep_low_resolution_observation(nat):
    """A high-resolution Observation + a LowResObservation of other bands (observation.py:242-599, row A14): the
    low-resolution planes go through scl_k_lowres (resample-and-convolve operator, residual, transpose), every
    iteration against the oracle's restated _render and its adjoint."""
    import os
    gold = np.load(os.path.join(helpers.GOLD, "ref_golden_lowres.npz"))
    frame, ohr, olr = helpers.lowres_scene(gold, "A_")
    srcs, blend = helpers.lowres_blend(gold, frame, ohr, olr)
    _lockstep(nat, blend, 10, e_rel=1e-9)


@pytest.mark.parametrize("approx", [True, False])
def test_scene_lockstep_low_resolution(nat, approx):
    """The multi-resolution scene forced through scene mode (tiled full-resolution planes + scl_k_lowres planes,
    losses and adjoint images combined), every iteration vs the oracle."""
    import os
    gold = np.load(os.path.join(helpers.GOLD, "ref_golden_lowres.npz"))
    frame, ohr, olr = helpers.lowres_scene(gold, "A_")
    srcs, blend = helpers.lowres_blend(gold, frame, ohr, olr)
    _lockstep(nat, blend, 6, e_rel=1e-9, approximate_L=approx, scene_mode=1)


@pytest.mark.parametrize("scene", [0, 1])
def test_low_resolution_planes_blocked_over_blends(nat, scene):
    """Blends that share a geometry: the low-resolution planes as matrix products over the blends (operator tiles
    reused by 64 blends, scl_k_lowres_gemm) against the per-blend kernels, and against the oracle."""
    import os
    gold = np.load(os.path.join(helpers.GOLD, "ref_golden_lowres.npz"))
    frame, ohr, olr = helpers.lowres_scene(gold, "A_")
    blends = []
    for i in range(3):
        srcs, bl = helpers.lowres_blend(gold, frame, ohr, olr)
        for c in bl.components:                    # three different states of the same scene
            c.sed = (c.sed * (1 + .05 * i)).astype(c.sed.dtype)
        blends.append(bl)
    packs = helpers.pack(blends)
    outs = []
    for flags in (4, 8):
        bt = nat.Batch(scene_mode=scene, **packs)
        bt.debug_flags(flags)
        bt.fit(6, 1e-9)
        outs.append(bt.download())
        bt.close()
    a, b = outs
    assert_allclose(b["logL"], a["logL"], rtol=5e-6)
    assert_array_equal(b["center"], a["center"])
    assert_allclose(b["morph"], a["morph"], rtol=0, atol=1e-5)
    for i, bl in enumerate(blends):
        for _ in range(6):
            bl.step(1e-9)
        assert_allclose(b["logL"][i, :6], bl.mse, rtol=1e-4)


def test_low_resolution_free_running_vs_reference(nat):
    """The same scene free-running against the reference's own 12 iterations (make_golden_lowres.py)."""
    import os
    gold = np.load(os.path.join(helpers.GOLD, "ref_golden_lowres.npz"))
    frame, ohr, olr = helpers.lowres_scene(gold, "A_")
    srcs, blend = helpers.lowres_blend(gold, frame, ohr, olr)
    ref = gold["A_mse"]
    bt = nat.Batch(**helpers.pack([blend]))
    bt.fit(len(ref), 1e-9)
    out = bt.download()
    bt.close()
    assert_allclose(out["logL"][0, :len(ref)], ref, rtol=1e-4)
    gm = gold["A_fit_model"]
    assert_allclose(_model(out, 0, len(srcs)), gm, rtol=0, atol=1e-4 * np.abs(gm).max())
    assert_array_equal(out["center"], gold["A_centers"])


# ---------------------------------------------------------------- scene mode (k_scene.cu)
def test_scene_mode_equals_fused_mode(nat):
    """The three-kernel K1 (render, overlap-save convolution tiles, adjoint) + HBM-scratch constraint kernels against
    the fused shared-memory path on the same blends (approximate_L, 20 iterations)."""
    blends = [helpers.synthetic_blend(s, kmin=3, kmax=6)[3] for s in (11, 12, 13)]
    outs = []
    for mode in (0, 1):
        bt = nat.Batch(scene_mode=mode, **helpers.pack(blends))
        bt.fit(20, 0.0, approximate_L=True)
        outs.append(bt.download())
        bt.close()
    a, b = outs
    assert_allclose(b["logL"], a["logL"], rtol=2e-5)         # 20 free-running fp32 iterations
    assert_array_equal(b["center"], a["center"])
    assert_allclose(b["sed"], a["sed"], rtol=0, atol=1e-5 * np.abs(a["sed"]).max())
    assert_allclose(b["morph"], a["morph"], rtol=0, atol=1e-5)


@pytest.mark.parametrize("approx", [True, False])
@pytest.mark.parametrize("tag", ["hsc", "sim"])
def test_scene_lockstep_npz(nat, tag, approx):
    """configs 1 and 2 forced through scene mode, with approximate_L (blend.py:189-202) and with the exact constants
    (Gram matrix by tiles, normalised squarings, Rayleigh quotient; blend.py:203-217), every iteration vs the oracle"""
    _, _, _, blend = helpers.npz_blend(tag)
    _lockstep(nat, blend, 12, approximate_L=approx, scene_mode=1)


@pytest.mark.parametrize("approx", [True, False])
def test_scene_lockstep_multicomponent(nat, approx):
    _, _, _, _, blend = helpers.multi_blend()
    _lockstep(nat, blend, 8, approximate_L=approx, scene_mode=1)


def test_scene_lockstep_large_frame(nat):
    """A 5-band 150x140 frame with 10-12 sources: beyond one convolution tile, beyond the shared-memory resident
    constraint kernels and beyond 8 components -- scene mode chosen automatically; every iteration vs the oracle."""
    _, _, srcs, blend = helpers.synthetic_blend(21, Ny=150, Nx=140, kmin=10, kmax=12)
    assert len(blend.components) > 8
    _lockstep(nat, blend, 4, approximate_L=True)
    _lockstep(nat, blend, 4, approximate_L=False)


def test_scene_lockstep_tall_frame_near_the_edge_limit(nat):
    """A 2-band 2040x48 frame, two pixels short of the 2042-pixel edge limit (include/scarlet_b200.h): the symmetry
    tables of the one-CTA constraint kernel pass 48 KB of dynamic shared memory there (the opt-in of ensure_globals), the
    level-ordered sweep weights exceed their budget (the sweep falls back to the offset-indexed table) and the
    offset tables hold 4079 x 95 entries.  Two lock-step iterations against the oracle through the chain of kernels,
    then one iteration of the one-CTA kernel (debug flag 16) against the chain."""
    _, _, _, blend = helpers.synthetic_blend(31, C=2, Ny=2040, Nx=48, kmin=2, kmax=2)
    p = helpers.pack([blend])
    outs = []
    for flags in (0, 16):
        bt = nat.Batch(**p)
        bt.debug_flags(flags)
        bt.fit(1, 1e-2)
        outs.append(bt.download())
        bt.close()
    a, b = outs
    assert_array_equal(a["center"], b["center"])
    assert_allclose(a["logL"], b["logL"], rtol=1e-6)
    assert_allclose(a["morph"], b["morph"], rtol=0, atol=2e-6)
    _lockstep(nat, blend, 2)


@pytest.mark.parametrize("approx", [False, True])
def test_scene_lockstep_config4_class_256(nat, approx):
    """BASELINE config 4's class at a quarter of its area: a 6-band 256x256 scene with 100 sources (70 %
    two-component => K ~ 170 > 64: several Gram tiles, the tiled squarings and the f64 Rayleigh quotient against
    np.linalg.eigvals of blend.py:203-217, ~70 shared-centre sources, the HBM level loop), three iterations vs the
    oracle: logL 1e-5, L 2e-6, centres / flags exact."""
    _, _, blend = helpers.scene_blend(256, 100)
    assert len(blend.components) > 128
    _lockstep(nat, blend, 3, approximate_L=approx)


def test_scene_lockstep_config4_full_512(nat):
    """ONE lock-step iteration of the real config-4 scene (6 bands, 512x512, 200 sources, K = 329) vs the oracle
    (~20 s of host time), exact Lipschitz constants."""
    _, _, blend = helpers.scene_blend(512, 200)
    assert len(blend.components) > 256
    _lockstep(nat, blend, 1)


@pytest.mark.parametrize("case", ["scene160", "blends64"])
def test_scene_early_exit_is_bit_identical(nat, case):
    """The scene-mode sweep stops at the first three consecutive levels without a positive pixel and the rest of
    the frame is zeroed (what the full sweep + prox_plus would leave there, operator.py:569-571, update.py:27-32):
    bit-identical state and loss history against the full sweep over several free-running iterations."""
    if case == "scene160":
        batch, _, _ = helpers.scene_blend(160, 24)
        kw = dict(batch)
    else:
        kw = dict(helpers.pack([helpers.synthetic_blend(s, kmin=3, kmax=6)[3] for s in (11, 12, 13)]), scene_mode=1)
    outs = []
    for flags in (0, 1):
        bt = nat.Batch(**kw)
        bt.debug_flags(flags)
        bt.fit(6, 0.0)
        outs.append(bt.download())
        bt.close()
    a, b = outs
    for key in ("sed", "morph", "center", "shift", "conv", "logL", "lipschitz"):
        assert_array_equal(a[key], b[key])
    if case == "scene160":
        assert (a["morph"] == 0).mean() > 0.2       # there was something to skip


@pytest.mark.parametrize("case", ["scene160", "ragged", "blends64"])
def test_scene_sweep_level_ordered_weights_are_bit_identical(nat, case):
    """The scene-mode sweep takes an interior pixel's weights from a level-ordered copy of the table that keeps the
    four taps that can be strictly closer to the peak (prox_dev.cuh: scl_k_build_wlev); debug flag 32 makes it read
    all eight from the offset-indexed table (operator.py:540-621).  Same products in the same order: identical bits,
    early exit or not, square or ragged frames."""
    if case == "scene160":
        kw = dict(helpers.scene_blend(160, 24)[0])
    elif case == "ragged":
        kw = dict(helpers.pack([helpers.synthetic_blend(21, Ny=150, Nx=140, kmin=10, kmax=12)[3]]))
    else:
        kw = dict(helpers.pack([helpers.synthetic_blend(s, kmin=3, kmax=6)[3] for s in (11, 12, 13)]), scene_mode=1)
    outs = []
    for flags in (0, 32, 33):
        bt = nat.Batch(**kw)
        bt.debug_flags(flags)
        bt.fit(6, 0.0, approximate_L=True)
        outs.append(bt.download())
        bt.close()
    for b in outs[1:]:
        for key in ("sed", "morph", "center", "shift", "conv", "logL", "lipschitz"):
            assert_array_equal(outs[0][key], b[key])


@pytest.mark.parametrize("case", ["scene160", "blends64", "multi"])
def test_scene_constraints_in_stages_equal_one_cta_kernel(nat, case):
    """Scene mode runs the constraint schedule as a chain of kernels (k_prox_split.cu: steps, source centres, symmetry
    tables, the two k-space products tiled over the GPU, the rest) -- against the one-CTA-per-component kernel
    (debug flag 16) over free-running iterations that include a centroid iteration (it % 5 == 0).  Everything but
    the k-space products is the same arithmetic; those sum in ascending l instead of the rotated order of half the
    warps, so a morphology may differ by float32 rounding of one product per iteration."""
    if case == "scene160":
        kw = dict(helpers.scene_blend(160, 24)[0])
    elif case == "multi":
        kw = dict(helpers.pack([helpers.multi_blend()[4]]), scene_mode=1)
    else:
        kw = dict(helpers.pack([helpers.synthetic_blend(s, kmin=3, kmax=6)[3] for s in (11, 12, 13)]), scene_mode=1)
    outs = []
    for flags in (0, 64, 16):
        bt = nat.Batch(**kw)
        bt.debug_flags(flags)
        bt.fit(7, 0.0)
        outs.append(bt.download())
        bt.close()
    a, c, b = outs
    # the chains on two streams (Gram beside the convolutions, the two groups of components beside each other) against
    # the same kernels in line on one stream: the same arithmetic
    # (not asserted bit for bit: the trace of a squaring is accumulated with atomics, in any order)
    for key in ("center", "conv"):
        assert_array_equal(a[key], c[key])
    for key in ("sed", "morph", "shift", "logL", "lipschitz"):
        assert_allclose(a[key], c[key], rtol=1e-6, atol=1e-9)
    assert_array_equal(a["center"], b["center"])
    assert_array_equal(a["conv"], b["conv"])
    assert_allclose(a["shift"], b["shift"], rtol=0, atol=1e-5)
    assert_allclose(a["logL"], b["logL"], rtol=5e-6)
    assert_allclose(a["lipschitz"], b["lipschitz"], rtol=1e-6)
    assert_allclose(a["morph"], b["morph"], rtol=0, atol=2e-6)
    assert_allclose(a["sed"], b["sed"], rtol=0, atol=2e-6 * np.abs(b["sed"]).max())


def test_scene_gram_by_footprint_boxes_equals_dense(nat):
    """Exact L in scene mode: S S^T over the intersections of the components' footprint boxes (scl_k_scene_bbox /
    scl_k_scene_gram_boxes) against the dense tiled contraction over every pixel of every pair: the same products,
    so the constants agree to float64 summation order and the iterates to a few ulp."""
    batch, _, _ = helpers.scene_blend(160, 24)
    outs = []
    for flags in (0, 2):
        bt = nat.Batch(**batch)
        bt.debug_flags(flags)
        bt.fit(4, 0.0)
        outs.append(bt.download())
        bt.close()
    a, b = outs
    assert_allclose(a["lipschitz"], b["lipschitz"], rtol=1e-6)
    assert_allclose(a["logL"], b["logL"], rtol=1e-6)
    assert_array_equal(a["center"], b["center"])
    assert_allclose(a["morph"], b["morph"], rtol=0, atol=2e-6)


def test_multicomponent_free_running_vs_reference(nat):
    """The same scene free-running against the reference's own run (tests/golden/make_golden_multi.py)."""
    g, _, _, srcs, blend = helpers.multi_blend()
    p = helpers.pack([blend])
    assert_allclose(p["morph"], g["init_morphs"], rtol=0, atol=2e-6)
    bt = nat.Batch(**p)
    bt.fit(int(g["max_iter"]), 1e-2)
    out = bt.download()
    bt.close()
    n = int(out["n_iter"][0])
    assert n == len(g["mse"])
    assert_allclose(out["logL"][0, :n], g["mse"], rtol=5 * REL)
    first = np.concatenate([[0], np.cumsum(g["n_comp"])])[:-1]
    assert_array_equal(out["center"][first], g["centers"])
    model = _model(out, 0, len(out["sed"]))
    assert_allclose(model, g["model"], rtol=0, atol=5 * REL * np.abs(g["model"]).max())


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_lockstep_synthetic(nat, seed):
    _, _, _, blend = helpers.synthetic_blend(seed)
    _lockstep(nat, blend, 12)


@pytest.mark.parametrize("C,Ny,Nx,psf", [(1, 39, 41, 15), (2, 38, 38, 11), (3, 47, 40, 21), (4, 41, 52, 15), (6, 40, 40, 21),
                                         (7, 39, 45, 11), (8, 52, 38, 21)])
def test_lockstep_band_counts_and_ragged_frames(nat, C, Ny, Nx, psf):
    """Every band count the kernels are instantiated for (1 .. 8), odd / non-square frames and small PSF stamps,
    every iteration vs the oracle (exact L and approximate L)."""
    _, _, srcs, blend = helpers.synthetic_blend(100 + C, C=C, Ny=Ny, Nx=Nx, kmin=2, kmax=3, psf_size=psf)
    assert len(srcs) >= 1
    _lockstep(nat, blend, 5)
    _lockstep(nat, blend, 3, approximate_L=True)


@pytest.mark.parametrize("tag", ["hsc", "sim"])
def test_free_running_vs_reference(nat, golden, tag):
    """Whole fit on the GPU vs the reference's own run (tests/golden): iteration count, per-iteration
    logL, centres and the final model."""
    _, _, _, blend = helpers.npz_blend(tag)
    bt = nat.Batch(**helpers.pack([blend]))
    bt.fit(200, 1e-2)
    out = bt.download()
    bt.close()
    ref = golden[tag + "_mse"]
    assert out["n_iter"][0] == len(ref)
    assert out["active"][0] == 0
    assert_allclose(out["logL"][0, :len(ref)], ref, rtol=5 * REL)
    assert_array_equal(out["center"], golden[tag + "_centers"])
    model = _model(out, 0, len(out["sed"]))
    assert_allclose(model, golden[tag + "_model"], rtol=0, atol=5 * REL * np.abs(golden[tag + "_model"]).max())


def test_batch_matches_single(nat):
    """A heterogeneous batch (different K) gives the same answers as blend-by-blend runs, and blends that
    converge early stop being updated."""
    blends = [helpers.synthetic_blend(s)[3] for s in range(6)]
    packs = helpers.pack(blends)
    bt = nat.Batch(**packs)
    info = bt.fit(40, 1e-2, check_every=4)
    out = bt.download()
    bt.close()
    assert info["launches"] > 0
    off = packs["comp_off"]
    for i, bl in enumerate(blends):
        b1 = nat.Batch(**helpers.pack([bl]))
        b1.fit(40, 1e-2)
        o1 = b1.download()
        b1.close()
        n = o1["n_iter"][0]
        assert out["n_iter"][i] == n
        assert_array_equal(out["logL"][i, :n], o1["logL"][0, :n])
        assert_array_equal(out["morph"][off[i]:off[i + 1]], o1["morph"])
        assert_array_equal(out["sed"][off[i]:off[i + 1]], o1["sed"])


def test_free_running_synthetic_vs_oracle(nat):
    blends = [helpers.synthetic_blend(s)[3] for s in range(4)]
    packs = helpers.pack(blends)
    bt = nat.Batch(**packs)
    bt.fit(25, 0.0)            # e_rel = 0: fixed iteration count
    out = bt.download()
    bt.close()
    off = packs["comp_off"]
    for i, bl in enumerate(blends):
        bl.fit(25, 0.0)
        assert out["n_iter"][i] == 25
        assert_allclose(out["logL"][i, :25], np.array(bl.mse, dtype=np.float64), rtol=20 * REL)
        assert_array_equal(out["center"][off[i]:off[i + 1]], [s.pixel_center for s in bl.sources])


def test_continue_fit_and_approximate_L(nat):
    """fit() called twice continues (blend.py:79, it = len(mse)); approximate_L path (blend.py:189-202)."""
    _, _, _, blend = helpers.npz_blend("hsc")
    ref = copy.deepcopy(blend)
    bt = nat.Batch(**helpers.pack([blend]))
    bt.fit(3, 1e-2, approximate_L=True)
    o = bt.download()
    bt.close()
    ref.fit(3, 1e-2, approximate_L=True)
    assert_allclose(o["logL"][0, :3], np.array(ref.mse, dtype=np.float64), rtol=REL)
    # second call starts from it0 = 3: centroid schedule keyed on the absolute iteration
    p = helpers.pack([ref])
    assert p["it0"][0] == 3
    bt = nat.Batch(**p)
    bt.fit(4, 1e-2)
    o = bt.download()
    bt.close()
    ref.fit(4, 1e-2)
    assert_allclose(o["logL"][0, :o["n_iter"][0]], np.array(ref.mse[3:], dtype=np.float64), rtol=REL)
    assert_array_equal(o["center"], [s.pixel_center for s in ref.sources])


def test_empty_and_errors(nat):
    _, _, _, blend = helpers.npz_blend("sim")
    p = helpers.pack([blend])
    bad = dict(p)
    bad["center"] = p["center"].copy()
    bad["center"][0] = (999, 0)
    with pytest.raises(ValueError):
        nat.Batch(**bad)
    bt = nat.Batch(**p)
    bt.fit(0)
    out = bt.download()
    assert out["n_iter"][0] == 0
    assert_array_equal(out["morph"], p["morph"].astype(np.float32))
    bt.close()


# ---------------------------------------------------------------- post-fit products (k_products.cu)
def _check_products(nat, blends, n_fit=6, **batch_kw):
    """Device products of a fitted batch against numpy on the downloaded state and the oracle's
    Observation.render(component.get_model()) (observation.py:198-220, component.py:148-175)."""
    from oracle import ref_blend as rb
    bt = nat.Batch(**helpers.pack(blends), **batch_kw)
    bt.fit(n_fit, 0.0, approximate_L=True)
    out = bt.download()
    flux, bounds, flags = bt.products(0.0)
    got = bt.render_components()
    part = bt.render_components(1, 2)
    bt.close()
    K = len(out["sed"])
    assert_array_equal(part, got[1:3])
    k = 0
    for bl in blends:
        obs = [o for o in bl.observations if not hasattr(o, "dense_operator")]
        for _ in bl.components:
            sed, morph = out["sed"][k], out["morph"][k]
            assert_allclose(flux[k], morph.sum() * sed, rtol=2e-6)
            nz = np.where(morph > 0)
            assert_array_equal(bounds[k], [nz[0].min(), nz[0].max(), nz[1].min(), nz[1].max()])
            edge = max(morph[:, 0].max(), morph[:, -1].max(), morph[0].max(), morph[-1].max()) > 0
            assert flags[k] == (out["conv"][k] & 3) | (4 if edge else 0)
            model = sed[:, None, None].astype(np.float64) * morph[None].astype(np.float64)
            want = np.concatenate([o.render(model) for o in obs])
            assert_allclose(got[k], want, rtol=0, atol=1e-5 * max(np.abs(want).max(), 1e-30))
            k += 1
    assert k == K
    return got, out


def test_products_fused_and_scene(nat):
    blends = [helpers.synthetic_blend(s, kmin=3, kmax=6)[3] for s in (11, 12, 13)]
    a, _ = _check_products(nat, blends)
    b, _ = _check_products(nat, blends, scene_mode=1)
    assert_allclose(b, a, rtol=0, atol=2e-5 * np.abs(a).max())


def test_products_large_frame_and_planes(nat):
    """Tiled renders on a 150x140 frame; several observations per blend render to their own planes."""
    _check_products(nat, [helpers.synthetic_blend(21, Ny=150, Nx=140, kmin=10, kmax=12)[3]], n_fit=3)
    _check_products(nat, [helpers.multi_obs_blend(5)[-1]], n_fit=4)


def test_products_flags_empty_and_edge(nat):
    """NO_VALID_PIXELS for an all-zero morphology, EDGE_PIXELS for flux on the border (BlendFlag, component.py:13-36)."""
    blend = helpers.synthetic_blend(3)[3]
    p = helpers.pack([blend])
    p["morph"] = np.array(p["morph"], dtype=np.float32)
    Ny, Nx = p["morph"].shape[-2:]
    p["morph"][0] = 0
    p["morph"][1] = 0
    p["morph"][1].reshape(Ny, Nx)[Ny - 1, 7] = 0.5
    bt = nat.Batch(**p)
    flux, bounds, flags = bt.products(0.0)
    bt.close()
    assert flags[0] & 8 and not flags[0] & 4 and list(bounds[0]) == [-1] * 4 and np.all(flux[0] == 0)
    assert flags[1] & 4 and not flags[1] & 8 and list(bounds[1]) == [Ny - 1, Ny - 1, 7, 7]
