"""SURVEY 8(f) row N4 inside the loop: custom update() schedules built from the constraints the reference exports
(update.py:13-103,106-199: positive_*, normalized, sparse_l0 / sparse_l1, threshold, monotonic, symmetric incl. the
sdss algorithm) and priors (component.py:39-67,177-187), every iteration against oracle sources that run the same
calls in Python (lock-step, north_star's tolerance), at the C-ABI level and through `import scarlet`."""
import numpy as np
import pytest
from numpy.testing import assert_array_equal, assert_allclose

import helpers
from oracle import ref_prox
from oracle import ref_blend as rb

pytestmark = pytest.mark.gpu
REL = 1e-5


@pytest.fixture(scope="module")
def nat():
    from scarlet_b200 import _native
    _native.require_device()
    return _native


# ---- oracle-side custom sources: the reference's functions called in the order the schedule lists them
def _step_morph(c):
    return 1 / c.L_morph


class SparseL1Source(rb.ExtendedSource):
    """update(): max_pixel, symmetric(sdss), sparse_l1(thresh), positive, normalized('morph_max')"""
    OPS = [("max_pixel", 0), ("sym_sdss", 0), ("sparse_l1", 0.02), ("positive_sed", 0), ("positive_morph", 0),
           ("norm_morph_max", 0)]

    def update(self):
        if not hasattr(self, "L_morph"):
            return rb.PointSource.update(self)                 # construction: the stock pass
        self.pixel_center = ref_prox.max_pixel(self.morph, self.pixel_center)
        ref_prox.prox_uncentered_symmetry(self.morph, 0, self.pixel_center, "sdss", None, self.shift, .5)
        ref_prox.prox_soft(self.morph, _step_morph(self), 0.02)
        ref_prox.prox_plus(self.sed)
        ref_prox.prox_plus(self.morph)
        ref_prox.normalized(self.sed, self.morph, "morph_max")
        return self


class MinimalSource(rb.ExtendedSource):
    """docs/user_docs.ipynb cell 38 (MinimallyConstrainedSource): positive_sed, positive_morph, normalized('sed')"""
    OPS = [("positive_sed", 0), ("positive_morph", 0), ("norm_sed", 0)]

    def update(self):
        if not hasattr(self, "L_morph"):
            return rb.PointSource.update(self)
        ref_prox.prox_plus(self.sed)
        ref_prox.prox_plus(self.morph)
        ref_prox.normalized(self.sed, self.morph, "sed")
        return self


class ThreshL0Source(rb.ExtendedSource):
    """max_pixel, centroid, symmetric(kspace), monotonic(thresh .05), sparse_l0, threshold, positive, normalized('morph')"""
    OPS = [("max_pixel", 0), ("centroid", 0), ("sym_kspace", 0), ("monotonic", .05), ("sparse_l0", 0.01), ("threshold", 0),
           ("positive_sed", 0), ("positive_morph", 0), ("norm_morph", 0)]

    def update(self):
        if not hasattr(self, "L_morph"):
            return rb.PointSource.update(self)
        self.pixel_center = ref_prox.max_pixel(self.morph, self.pixel_center)
        self.pixel_center, self.shift = ref_prox.psf_weighted_centroid(self.morph, self.centroid_weight, self.pixel_center)
        ref_prox.prox_uncentered_symmetry(self.morph, 0, self.pixel_center, "kspace", None, self.shift, .5)
        m = self.morph.copy()
        ref_prox.build_monotonic(m.shape, self.pixel_center, .05)(m)
        self.morph[:] = m
        ref_prox.prox_hard(self.morph, _step_morph(self), 0.01)
        ref_prox.update_threshold(self.morph)
        ref_prox.prox_plus(self.sed)
        ref_prox.prox_plus(self.morph)
        ref_prox.normalized(self.sed, self.morph, "morph")
        return self


def _blend(seed, classes, priors=None):
    """A config-3 style blend whose sources cycle through `classes` (stock ExtendedSource where None)."""
    frame, obs, srcs, _ = helpers.synthetic_blend(seed, kmin=3, kmax=5)
    out = []
    for i, s in enumerate(srcs):
        cls = classes[i % len(classes)]
        if cls is not None:
            s.__class__ = cls
        if priors is not None and priors[i % len(priors)] is not None:
            s.prior = priors[i % len(priors)](s)
        out.append(s)
    return rb.Blend(out, obs)


def _pack(nat, blend):
    from scarlet_b200 import schedule
    p = helpers.pack([blend])
    K = len(blend.components)
    ops, args = np.zeros((K, nat.SCHED_OPS), np.int32), np.zeros((K, nat.SCHED_OPS), np.float32)
    any_s = False
    par = np.tile(np.float32([0, 0, -1, -1]), (K, 1))
    b_sed, idx, b_morph = np.zeros((K, blend.frame.shape[0]), np.float32), -np.ones(K, np.int32), []
    any_p = False
    for k, s in enumerate(blend.sources):
        if hasattr(s, "OPS"):
            any_s = True
            p["cflags"][k] = nat.F_SCHEDULE
            for i, (name, arg) in enumerate(s.OPS):
                ops[k, i], args[k, i] = schedule.OPS[name], arg
        pr = getattr(s, "prior", None)
        if pr is not None:
            any_p = True
            par[k] = pr.par
            b_sed[k] = pr.b_sed
            if pr.b_morph is not None:
                idx[k] = len(b_morph)
                b_morph.append(pr.b_morph)
    if any_s:
        p["sched"] = dict(ops=ops, args=args)
    if any_p:
        p["prior"] = dict(par=par, b_sed=b_sed, idx=idx, b_morph=np.array(b_morph, dtype=np.float32))
    return p


def _lockstep(nat, blend, n_steps, approximate_L=False):
    for step in range(n_steps):
        bt = nat.Batch(**_pack(nat, blend))
        bt.fit(1, 1e-2, approximate_L=approximate_L)
        out = bt.download()
        bt.close()
        conv = blend.step(1e-2, approximate_L=approximate_L)
        assert_allclose(out["logL"][0, 0], blend.mse[-1], rtol=REL)
        assert_array_equal(out["center"], [s.pixel_center for s in blend.sources])
        assert_allclose(out["shift"], [s.shift for s in blend.sources], rtol=0, atol=5e-6)
        sed = np.array([c.sed for c in blend.components])
        morph = np.array([c.morph for c in blend.components])
        assert_allclose(out["sed"], sed, rtol=REL, atol=REL * np.abs(sed).max())
        scale = np.abs(morph).max(axis=(1, 2))[:, None, None] + 1e-30          # 'sed' / 'morph' normalisations rescale it
        assert_allclose(out["morph"] / scale, morph / scale, rtol=REL, atol=REL)
        assert_array_equal(out["conv"], [c.flags for c in blend.components])
        assert out["active"][0] == (0 if conv else 1)
        if conv:
            break


@pytest.mark.parametrize("classes", [[SparseL1Source, None], [MinimalSource], [ThreshL0Source, None, SparseL1Source]])
def test_lockstep_declared_schedules(nat, classes):
    blend = _blend(43, classes)       # a seed on which no SED is driven to zero (normalized('sed') divides by its sum)
    _lockstep(nat, blend, 8)


class _AffineOracle(rb.Prior):
    """The oracle side of an affine prior: plain callbacks, as the reference's Prior takes them."""

    def __init__(self, a_sed, b_sed, a_morph, b_morph):
        self.par = (a_sed, a_morph, abs(a_sed), abs(a_morph))
        self.b_sed, self.b_morph = b_sed, b_morph
        bm = 0 if b_morph is None else b_morph
        rb.Prior.__init__(self, lambda sed, morph: (a_sed * sed + b_sed, a_morph * morph + bm),
                          lambda sed, morph: (abs(a_sed), abs(a_morph)))


def test_lockstep_priors(nat):
    """A Gaussian prior pulling the morphology towards its initial template and damping the SED, on stock sources,
    a linear morphology penalty on a source with a declared schedule, one source without a prior; exact and
    approximate Lipschitz constants."""
    def gauss(s):
        lam = 0.3 * float(np.sum(s.sed ** 2))
        return _AffineOracle(1e-3, np.zeros(len(s.sed), np.float32), lam, (-lam * s.morph).astype(np.float32))

    def linear(s):
        return _AffineOracle(0.0, np.full(len(s.sed), 0.5, np.float32), 0.0, np.full(s.morph.shape, 2.0, np.float32))
    for approx in (False, True):
        blend = _blend(42, [None, SparseL1Source, None], priors=[gauss, linear, None])
        _lockstep(nat, blend, 6, approximate_L=approx)


def test_fix_flags_skip_the_prior(nat):
    """component.py:181-186: a fixed parameter takes neither the prior's gradient nor its Lipschitz constant."""
    blend = _blend(42, [None], priors=[lambda s: _AffineOracle(2e-3, np.zeros(len(s.sed), np.float32), 5.0, None)])
    blend.components[0].fix_morph = True
    blend.components[1].fix_sed = True
    _lockstep(nat, blend, 4)


# ---- through the Python surface: update() overrides are traced
def test_traced_update_through_import_scarlet(nat):
    import scarlet
    from scarlet import update
    frame_o, obs_o, srcs_o, _ = helpers.synthetic_blend(43, kmin=3, kmax=5)

    class Minimal(scarlet.ExtendedSource):
        def update(self):
            update.positive_sed(self)
            update.positive_morph(self)
            update.normalized(self, 'sed')
            return self

    class Sparse(scarlet.ExtendedSource):
        def update(self):
            self.pixel_center = scarlet.measurement.max_pixel(self.morph, self.pixel_center)
            update.symmetric(self, self.pixel_center, algorithm="sdss")
            update.sparse_l1(self, 0.02)
            update.positive(self)
            update.normalized(self, 'morph_max')
            return self

    class IterationDependent(scarlet.ExtendedSource):
        n = 0

        def update(self):
            IterationDependent.n += 1
            if IterationDependent.n % 2:
                update.positive(self)
            return self

    mpsf = scarlet.psf.generate_psf_image(scarlet.psf.gaussian, (41, 41), sigma=.9).image
    frame = scarlet.Frame(obs_o.images.shape, psfs=mpsf[None])
    obs = scarlet.Observation(obs_o.images, psfs=obs_o.frame.psfs, weights=obs_o.weights).match(frame)
    bg = np.ones(5, np.float32)
    classes, oclasses = [Sparse, Minimal, scarlet.ExtendedSource], [SparseL1Source, MinimalSource, None]
    srcs = []
    for i, so in enumerate(srcs_o):
        # start both sides from the oracle's stock initial state (the custom update() of the product classes ran
        # once at construction; overwrite with the common state)
        s = classes[i % 3](frame, so.pixel_center_init, obs, bg)
        s._sed, s._morph = so.sed.copy(), so.morph.copy()
        s.pixel_center, s.shift = so.pixel_center, so.shift
        if oclasses[i % 3] is not None:
            so.__class__ = oclasses[i % 3]
        srcs.append(s)
    blend = scarlet.Blend(srcs, obs)
    ref = rb.Blend(srcs_o, obs_o)
    blend.fit(6, e_rel=0)
    for _ in range(6):
        ref.step(0)
    assert blend.it == 6
    assert_allclose(blend.mse, np.array(ref.mse, dtype=np.float64), rtol=20 * REL)
    for s, so in zip(blend.sources, ref.sources):
        assert s.pixel_center == so.pixel_center
        assert_allclose(s.morph, so.morph, rtol=0, atol=2e-4 * np.abs(so.morph).max())
    bad = IterationDependent(frame, srcs_o[0].pixel_center_init, obs, bg)
    with pytest.raises(NotImplementedError):
        scarlet.Blend([bad], obs).fit(1)
    with pytest.raises(NotImplementedError):
        c = scarlet.Component(frame, np.ones(5), np.ones((64, 64)), prior=scarlet.Prior(lambda a, b: (0, 0), lambda a, b: (0, 0)))
        scarlet.Blend([c], obs).fit(1)
    pc = scarlet.Component(frame, np.ones(5), np.ones((64, 64)) * .01, prior=scarlet.AffinePrior(a_morph=.1))
    scarlet.Blend([pc] + srcs[:1], obs).fit(2)
