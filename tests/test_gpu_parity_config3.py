"""north_star's tolerance on the headline workload (BASELINE config 3: 5-band 64x64 blends, 2-8 ExtendedSources,
fit(200, 1e-2)): the free-running device fit is re-seeded into the oracle at iterations 30, 100 and one before the
last, and ONE iteration from that common state must agree at rel 1e-5 (logL, SEDs, morphologies; centres and flags
exact) -- the trajectory of an fp32 descent may drift, a single iteration may not.  Plus: on how many blends the
device and the oracle stop on a different iteration."""
import copy

import numpy as np
import pytest
from numpy.testing import assert_array_equal, assert_allclose

import helpers
from test_gpu_fit import _lockstep, _model

pytestmark = pytest.mark.gpu

N_BLENDS = 64


@pytest.fixture(scope="module")
def nat():
    from scarlet_b200 import _native
    _native.require_device()
    return _native


@pytest.fixture(scope="module")
def blends():
    return [helpers.synthetic_blend(300 + s)[3] for s in range(N_BLENDS)]


def _run(nat, packs, n):
    bt = nat.Batch(**packs)
    bt.fit(n, 1e-2)
    out = bt.download()
    bt.close()
    return out


def test_reseeded_single_iterations_and_stopping_counts(nat, blends):
    packs = helpers.pack(blends)
    off = packs["comp_off"]
    full = _run(nat, packs, 200)
    n_dev = full["n_iter"].copy()
    assert (full["active"] == 0).mean() > 0.9
    checked = 0
    for T in (30, 100):
        out = _run(nat, packs, T)
        for i, bl in enumerate(blends):
            if out["n_iter"][i] != T or not out["active"][i]:
                continue                                       # stopped before T
            assert_array_equal(out["logL"][i, :T], full["logL"][i, :T])     # the same trajectory as the full run
            seeded = copy.deepcopy(bl)
            helpers.seed_oracle(seeded, out, off[i], out["logL"][i, :T])
            _lockstep(nat, seeded, 1)
            checked += 1
    # one iteration before the device's last: the common step must also agree on "converged now"
    for i in range(0, N_BLENDS, 8):
        n = int(n_dev[i])
        if n < 3 or full["active"][i]:
            continue
        p1 = helpers.pack([blends[i]])
        out = _run(nat, p1, n - 1)
        assert out["active"][0] == 1
        seeded = copy.deepcopy(blends[i])
        helpers.seed_oracle(seeded, out, 0, out["logL"][0, :n - 1])
        _lockstep(nat, seeded, 1)
        assert seeded.converged
        checked += 1
    assert checked >= 20
    # free-running oracle: stopping iteration, loss history and final model.  63 of these 64 blends end within ~1e-6 of
    # the oracle (tools/parity_probe.py --converge 64).  A trajectory can also hit one of the reference's own
    # discontinuities -- the `x <= 0` mask of prox_kspace_symmetry (operator.py:270,287), an argmax, a rounded centroid --
    # where rounding decides the branch: there the reference's arithmetic run in float32 and in float64 part ways as
    # well (seed 355: 5e-3 on a morphology from iteration 2 on), and the device run must then agree with one of the two.
    differ, n_f64, worst_model, worst_logl = 0, 0, 0.0, 0.0
    for i, bl in enumerate(blends):
        ref = copy.deepcopy(bl)
        ref.fit(200, 1e-2)
        if ref.it != n_dev[i]:
            differ += 1
            continue
        got = _model(full, off[i], off[i + 1])

        def deviation(r):
            n = r.it
            want = r.get_model().astype(np.float64)
            return (np.max(np.abs(full["logL"][i, :n] / np.array(r.mse, dtype=np.float64) - 1)),
                    np.abs(got - want).max() / np.abs(want).max())
        dl, dm = deviation(ref)
        if dm > 2e-5 or dl > 1e-4:
            ref64 = helpers.synthetic_blend(300 + i, dtype=np.float64)[3]
            ref64.fit(200, 1e-2)
            assert ref64.it == n_dev[i]
            dl, dm = deviation(ref64)
            n_f64 += 1
        worst_logl, worst_model = max(worst_logl, dl), max(worst_model, dm)
    print("config 3, %d blends: %d stop on a different iteration than the oracle; %d follow the float64 branch of a "
          "discontinuity; worst free-running deviation logL %.2e, model %.2e of the peak"
          % (N_BLENDS, differ, n_f64, worst_logl, worst_model))
    assert differ <= 0.02 * N_BLENDS + 1e-9 or differ <= 1
    assert n_f64 <= 0.05 * N_BLENDS
    assert worst_logl < 1e-4 and worst_model < 2e-5
