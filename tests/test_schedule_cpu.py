"""Host logic of the declared schedules and affine priors (scarlet_b200/schedule.py, component.AffinePrior,
blend.pack_blends): tracing a custom update() into op codes, refusing what cannot be traced, packing the descriptor
arrays.  No device needed."""
import numpy as np
import pytest

import scarlet
from scarlet import update
from scarlet_b200 import schedule, _native
from scarlet_b200.blend import pack_blends


def _frame_obs(C=3, Ny=12, Nx=10):
    frame = scarlet.Frame((C, Ny, Nx), psfs=None)
    obs = scarlet.Observation(np.zeros((C, Ny, Nx), np.float32)).match(frame)
    return frame, obs


class Minimal(scarlet.Component):
    """docs/user_docs.ipynb cell 38"""
    def update(self):
        update.positive_sed(self)
        update.positive_morph(self)
        update.normalized(self, 'sed')
        return self


class Monotonic(scarlet.Component):
    """docs/user_docs.ipynb cell 41, with the functions this version of the reference has"""
    pixel_center = (6, 5)

    def update(self):
        self.pixel_center = scarlet.measurement.max_pixel(self.morph, self.pixel_center)
        update.symmetric(self, self.pixel_center, algorithm="soft", strength=.25)
        update.monotonic(self, self.pixel_center, thresh=.1)
        update.sparse_l1(self, 0.5)
        update.threshold(self)
        update.positive(self)
        update.normalized(self, 'morph_max')
        return self


def test_trace_lists_the_calls_in_order():
    frame, obs = _frame_obs()
    m = Minimal(frame, np.ones(3), np.ones((12, 10)))
    O = schedule.OPS
    assert schedule.is_custom(m)
    assert schedule.trace(m) == [(O["positive_sed"], 0.0), (O["positive_morph"], 0.0), (O["norm_sed"], 0.0)]
    s = Monotonic(frame, np.ones(3), np.ones((12, 10)))
    ops = schedule.trace(s)
    assert [o for o, _ in ops] == [O["max_pixel"], O["sym_soft"], O["monotonic"], O["sparse_l1"], O["threshold"],
                                   O["positive_sed"], O["positive_morph"], O["norm_morph_max"]]
    assert [a for _, a in ops][1:4] == [0.25, pytest.approx(0.1), 0.5]
    assert s.pixel_center == (6, 5) and type(s.pixel_center) is tuple          # the live-centre marker does not leak
    assert not schedule.is_custom(scarlet.Component(frame, np.ones(3), np.ones((12, 10))))
    # nothing was computed while tracing
    assert np.all(s.morph == 1) and np.all(s.sed == 1)


def test_untraceable_updates_are_refused():
    frame, obs = _frame_obs()

    class Flip(scarlet.Component):
        n = 0

        def update(self):
            Flip.n += 1
            if Flip.n % 2:
                update.positive(self)
            return self

    class FixedCentre(scarlet.Component):
        pixel_center = (6, 5)

        def update(self):
            update.monotonic(self, (2, 2))
            return self

    class Nearest(scarlet.Component):
        pixel_center = (6, 5)

        def update(self):
            update.monotonic(self, self.pixel_center, use_nearest=True)
            return self

    class Shift(scarlet.Component):
        shift = (0.1, 0.2)

        def update(self):
            update.translation(self)
            return self

    for cls in (Flip, FixedCentre, Nearest, Shift):
        with pytest.raises(NotImplementedError):
            schedule.trace(cls(frame, np.ones(3), np.ones((12, 10))))


def test_affine_prior_is_a_reference_style_prior():
    frame, obs = _frame_obs()
    b = np.arange(120, dtype=np.float32).reshape(12, 10)
    p = scarlet.AffinePrior(a_sed=0.5, b_sed=[1, 2, 3], a_morph=-2.0, b_morph=b)
    c = scarlet.Component(frame, np.ones(3), np.full((12, 10), 2.0), prior=p)
    p.compute_grad(c)                                    # component.py:58-67
    np.testing.assert_allclose(p.sed_grad, 0.5 + np.array([1, 2, 3]))
    np.testing.assert_allclose(p.morph_grad, -4.0 + b)
    assert (p.L_sed, p.L_morph) == (0.5, 2.0)


def test_pack_blends_ships_schedules_and_priors():
    frame, obs = _frame_obs()
    plain = scarlet.Component(frame, np.ones(3), np.ones((12, 10)))
    mini = Minimal(frame, np.ones(3), np.ones((12, 10)), fix_sed=True)
    pri = scarlet.Component(frame, np.ones(3), np.ones((12, 10)), prior=scarlet.AffinePrior(a_morph=3.0, b_sed=1.0))
    packed = pack_blends([scarlet.Blend([plain, mini, pri], obs)])
    assert list(packed["cflags"]) == [0, _native.F_SCHEDULE | _native.F_FIX_SED, 0]
    assert packed["sched"]["ops"].shape == (3, _native.SCHED_OPS)
    assert list(packed["sched"]["ops"][1][:4]) == [1, 2, 3, 0] and not packed["sched"]["ops"][[0, 2]].any()
    par = packed["prior"]["par"]
    assert par[0, 2] < 0 and par[1, 2] < 0                      # no prior
    assert list(par[2]) == [0.0, 3.0, 0.0, 3.0]
    assert list(packed["prior"]["b_sed"][2]) == [1, 1, 1] and packed["prior"]["idx"][2] == -1
    cb = scarlet.Component(frame, np.ones(3), np.ones((12, 10)), prior=scarlet.Prior(lambda s, m: (0, 0), lambda s, m: (0, 0)))
    with pytest.raises(NotImplementedError):
        pack_blends([scarlet.Blend([cb], obs)])
