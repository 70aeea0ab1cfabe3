"""Declared schedules: how a custom `update()` runs inside the CUDA loop.

The reference lets a source class override `Component.update` with any Python code built from `scarlet.update.*`
(docs/user_docs.ipynb cells 35-41); `Blend.fit` calls it every iteration (component.py:359-367).  A Python callback
cannot run inside the device loop, so the override is TRACED once on the host: while `trace()` runs the method, the
functions of `scarlet.update` and the two centre measurements of `scarlet.measurement` record (op code, parameter)
instead of computing, and the resulting list -- at most SCHED_OPS entries -- is what the constraint kernel interprets
for that component every iteration (csrc/k_prox.cu: scl_run_schedule).  What can be traced is therefore exactly what
those modules export; a method whose op list depends on the iteration number, or that computes on `morph` / `sed` by
other means, raises NotImplementedError instead of fitting something else.
"""
import threading

from . import _native

# op codes (include/scarlet_b200.h: SCL_OP_*)
OPS = dict(positive_sed=1, positive_morph=2, norm_sed=3, norm_morph=4, norm_morph_max=5, sparse_l0=6, sparse_l1=7,
           threshold=8, monotonic=9, sym_kspace=10, sym_soft=11, sym_sdss=12, max_pixel=13, centroid=14)

_state = threading.local()


class LiveCenter(tuple):
    """`pixel_center` as the device tracks it: what measurement.max_pixel / psf_weighted_centroid return while a
    schedule is being traced."""


class LiveShift(tuple):
    """`shift` as the device tracks it."""


def recording():
    """The op list being recorded by the current thread, or None."""
    return getattr(_state, "rec", None)


def record(op, arg=0.0):
    rec = recording()
    if len(rec) >= _native.SCHED_OPS:
        raise NotImplementedError("an update() schedule holds at most %d operations" % _native.SCHED_OPS)
    rec.append((OPS[op], float(arg)))


def live_center(component, pixel_center, what):
    """The device applies centre-dependent operators at the component's own (live) pixel_center."""
    own = getattr(component, "pixel_center", None)
    if isinstance(pixel_center, LiveCenter) or (own is not None and tuple(pixel_center) == tuple(own)):
        return
    raise NotImplementedError("%s inside the CUDA loop works at the component's own pixel_center" % what)


def _trace_once(source):
    saved = {k: getattr(source, k) for k in ("pixel_center", "shift") if hasattr(source, k)}
    _state.rec = []
    try:
        source.update()
        return list(_state.rec)
    finally:
        _state.rec = None
        for k, v in saved.items():
            setattr(source, k, v)


def trace(source):
    """Op list [(code, arg), ...] of `source.update()`.  Traced twice: the list must not change between calls."""
    first = _trace_once(source)
    if _trace_once(source) != first:
        raise NotImplementedError("update() makes different calls from one iteration to the next; "
                                  "only fixed schedules run inside the CUDA loop")
    return first


def is_custom(source):
    """True when the source's update() is user code rather than one of the built-in schedules."""
    from .component import Component, ComponentTree
    from .source import PointSource, MultiComponentSource
    if isinstance(source, MultiComponentSource):
        return False
    if isinstance(source, ComponentTree):
        return False
    if isinstance(source, PointSource):
        return type(source).update is not PointSource.update
    return type(source).update is not Component.update
