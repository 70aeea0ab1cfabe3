"""Blend: the scene and its fit (mirrors scarlet/blend.py).

`Blend.fit` packs the components and observations into device buffers, runs the
whole proximal-gradient loop (blend.py:65-102 of the reference) as CUDA kernels
through the C ABI, and writes SEDs, morphologies, centres, flags and the loss
history back.  `fit_many` does the same for any number of independent blends in
one batched launch sequence (the reference has no batch notion; Blend.fit is a
batch of one).
"""
import logging

import numpy as np

from .component import ComponentTree, Component, BlendFlag, AffinePrior
from . import schedule
from .observation import LowResObservation
from . import _native
from .source import PointSource, MultiComponentSource, component_flags, _centroid_weight

logger = logging.getLogger("scarlet.blend")


def _owner_of(blend):
    """(source, component) pairs in component order."""
    pairs = []
    for s in blend.sources:
        comps = s.components if isinstance(s, ComponentTree) else [s]
        for c in comps:
            if not any(c is p[1] for p in pairs):
                pairs.append((s, c))
    return pairs


def pack_blends(blends, want_spectra=True, dtype=np.float32):
    """Arrays for _native.Batch from Blend objects that share one frame shape."""
    C, Ny, Nx = blends[0].frame.shape
    comp_off, sed, morph, cflags, center, shift, conv, src_first = [0], [], [], [], [], [], [], []
    images, weights, specs, ker_idx, cws, cw_idx, it0, prev = [], [], [], [], [], [], [], []
    spec_ids, cw_ids, lr_ids = {}, {}, {}
    sched_rows, prior_rows = {}, {}            # component index -> (ops, args) / AffinePrior
    lr_images, lr_weights, lr_ops, lr_op_idx = [], [], [], []
    any_w = any(isinstance(o.weights, np.ndarray) for b in blends for o in b.observations)
    # image planes: the bands of every observation stacked, each looking at one model band (observation.py:184-189)
    plane_band = None
    for bl in blends:
        if bl.frame.shape != (C, Ny, Nx):
            raise ValueError("all blends of a batch must share the frame shape")
        if bl.frame.dtype != dtype:
            raise NotImplementedError("batches compute in float32 (float64 blends go through fit_f64, one at a time)")
        pb, pb_lr = [], []
        for obs in bl.observations:
            if isinstance(obs, LowResObservation):
                if obs._lr_op is None:
                    raise ValueError("LowResObservation.match(frame) has not been called")
                pb_lr += list(range(C))[obs._band_slice]
                continue
            if obs.images.shape[1:] != (Ny, Nx):
                raise ValueError("observation and model frame differ in shape")
            sl = obs._band_slice
            pb += list(range(C))[sl]
            if len(list(range(C))[sl]) != obs.images.shape[0]:
                raise ValueError("observation bands do not match its slice of the model channels")
        n_lr_planes = len(pb_lr)
        if n_lr_planes and len([o for o in bl.observations if isinstance(o, LowResObservation)]) > 1:
            raise NotImplementedError("one LowResObservation per blend")
        pb = pb + pb_lr                      # full-resolution planes first, low-resolution planes after
        if plane_band is None:
            plane_band = pb
        elif pb != plane_band:
            raise ValueError("all blends of a batch must share the layout of their observations")
        if len(pb) > 8:
            raise NotImplementedError("at most 8 image planes (bands summed over observations) per blend")
        last_src = None
        for s, c in _owner_of(bl):
            if c.prior is not None and not isinstance(c.prior, AffinePrior):
                raise NotImplementedError("Prior callbacks cannot run inside the CUDA loop (AffinePrior can)")
            if isinstance(s, ComponentTree) and not isinstance(s, MultiComponentSource):
                raise NotImplementedError("only MultiComponentSource trees can run inside the CUDA loop")
            if isinstance(s, MultiComponentSource):
                if type(s).update is not MultiComponentSource.update or type(c).update is not Component.update:
                    raise NotImplementedError("custom update() overrides of multi-component sources cannot run "
                                              "inside the CUDA loop")
            if schedule.is_custom(s):
                # a custom update(): traced once into the op list the constraint kernel interprets (schedule.py)
                ops = schedule.trace(s)
                row_o, row_a = np.zeros(_native.SCHED_OPS, np.int32), np.zeros(_native.SCHED_OPS, np.float32)
                for i, (code, arg) in enumerate(ops):
                    row_o[i], row_a[i] = code, arg
                sched_rows[len(sed)] = (row_o, row_a)
                cflags.append(_native.F_SCHEDULE | (_native.F_FIX_SED if c.fix_sed else 0) |
                              (_native.F_FIX_MORPH if c.fix_morph else 0))
            else:
                cflags.append(component_flags(s, c))
            if c.prior is not None:
                prior_rows[len(sed)] = c.prior
            if not (isinstance(s, ComponentTree) and len(sed) > comp_off[-1] and last_src is s):
                first = len(sed)
            last_src = s
            src_first.append(first)
            sed.append(c._sed)
            morph.append(c._morph)
            center.append(getattr(s, "pixel_center", (Ny // 2, Nx // 2)))
            shift.append(getattr(s, "shift", (0., 0.)))
            conv.append(c.flags.value & 3)
        comp_off.append(len(sed))
        for obs in bl.observations:
            if isinstance(obs, LowResObservation):
                ys, xs = obs._window()
                sub = obs.images[:, ys, xs]
                lr_images.append(sub.reshape(len(sub), -1))
                w = obs.weights if isinstance(obs.weights, np.ndarray) else np.float32(obs.weights)
                lr_weights.append(np.broadcast_to(w, sub.shape).reshape(len(sub), -1))
                key = id(obs._lr_op)
                if key not in lr_ids:
                    lr_ids[key] = len(lr_ops)
                    lr_ops.append(obs._lr_op)
                lr_op_idx.append(lr_ids[key])
                continue
            Co = obs.images.shape[0]
            images.extend(obs.images)
            if any_w:
                weights.extend(obs.weights if isinstance(obs.weights, np.ndarray)
                               else np.ones(obs.images.shape, np.float32) * obs.weights)
            spec = obs.kernel_spectrum() if want_spectra else None
            if spec is None:
                ker_idx += [-1] * Co
            else:
                key = id(obs._diff_kernels.image) if len(blends) > 64 else obs._diff_kernels.image.tobytes()
                if key not in spec_ids:
                    spec_ids[key] = sum(len(sp) for sp in specs)
                    specs.append(spec)
                ker_idx += [spec_ids[key] + c for c in range(Co)]
        cw = np.asarray(_centroid_weight(bl.frame), dtype=np.float32)
        key = id(bl.frame.psfs) if len(blends) > 64 else cw.tobytes()
        if key not in cw_ids:
            cw_ids[key] = len(cws)
            cws.append(cw)
        cw_idx.append(cw_ids[key])
        it0.append(bl.it)
        prev.append(float(bl.mse[-1]) if bl.mse else 0.)
    K = len(sed)
    sched = prior = None
    if sched_rows:
        sched = dict(ops=np.zeros((K, _native.SCHED_OPS), np.int32), args=np.zeros((K, _native.SCHED_OPS), np.float32))
        for k, (o, a) in sched_rows.items():
            sched["ops"][k], sched["args"][k] = o, a
    if prior_rows:
        prior = dict(par=np.tile(np.float32([0, 0, -1, -1]), (K, 1)), b_sed=np.zeros((K, C), np.float32),
                     idx=-np.ones(K, np.int32), b_morph=[])
        for k, pr in prior_rows.items():
            prior["par"][k] = (pr.a_sed, pr.a_morph, pr._L[0], pr._L[1])
            prior["b_sed"][k] = np.broadcast_to(np.asarray(pr.b_sed, dtype=np.float32), (C,))
            if np.any(np.asarray(pr.b_morph) != 0):
                prior["idx"][k] = len(prior["b_morph"])
                prior["b_morph"].append(np.broadcast_to(np.asarray(pr.b_morph, dtype=np.float32), (Ny, Nx)))
        prior["b_morph"] = np.array(prior["b_morph"], dtype=np.float32).reshape(-1, Ny * Nx)
    return dict(C=C, Ny=Ny, Nx=Nx, comp_off=np.array(comp_off), sed=np.array(sed), morph=np.array(morph), sched=sched,
                prior=prior,
                cflags=np.array(cflags), center=np.array(center), shift=np.array(shift, dtype=np.float64),
                conv=np.array(conv), images=np.array(images), weights=np.array(weights) if any_w else None,
                kernel_spec=np.concatenate(specs) if specs else None, ker_idx=np.array(ker_idx),
                cw=np.array(cws), cw_idx=np.array(cw_idx), it0=np.array(it0), prev_loss=np.array(prev),
                src_first=np.array(src_first), plane_band=np.array(plane_band, dtype=np.int32),
                n_obs=len(blends[0].observations),
                ker_half=max([_native.kernel_half(*o._diff_kernels.image.shape[1:], Ny, Nx) for b in blends for o in b.observations
                              if not isinstance(o, LowResObservation) and o._diff_kernels is not None] + [0]),
                lr=dict(images=np.array(lr_images), weights=np.array(lr_weights), op=np.array(lr_ops),
                        op_idx=np.array(lr_op_idx)) if lr_images else None)


def unpack_blends(blends, packed, out):
    """Write the device results back into the Python objects."""
    off = packed["comp_off"]
    for i, bl in enumerate(blends):
        n = int(out["n_iter"][i])
        if out["logL"] is not None:
            bl.mse.extend(float(v) for v in out["logL"][i, :n])
        if n > 0:
            bl.L_sed, bl.L_morph = float(out["lipschitz"][i, 0]), float(out["lipschitz"][i, 1])
        for j, (s, c) in enumerate(_owner_of(bl)):
            k = off[i] + j
            c._sed = out["sed"][k].copy()
            c._morph = out["morph"][k].copy()
            if n > 0:
                c._last_sed, c._last_morph = c._sed.copy(), c._morph.copy()      # blend.py:180-182
            c.flags = (c.flags & ~(BlendFlag.SED_NOT_CONVERGED | BlendFlag.MORPH_NOT_CONVERGED)) | BlendFlag(int(out["conv"][k]))
            c.L_sed, c.L_morph = getattr(bl, "L_sed", 1), getattr(bl, "L_morph", 1)
            if isinstance(s, (PointSource, MultiComponentSource)) or (schedule.is_custom(s) and hasattr(s, "pixel_center")):
                s.pixel_center = (int(out["center"][k, 0]), int(out["center"][k, 1]))
                if getattr(s, "symmetric", False) or (schedule.is_custom(s) and hasattr(s, "shift")):
                    s.shift = (float(out["shift"][k, 0]), float(out["shift"][k, 1]))


def fit_f64(bl, max_iter=200, e_rel=1e-2, approximate_L=False):
    """`Blend.fit` of one blend whose frame is float64 (the reference's own loop tests, tests/test_blend.py:56-122):
    the same packing as the float32 batches, run by the float64 kernels (scl_fit_f64, csrc/k_f64.cu)."""
    fr = bl.frame
    if any(isinstance(o, LowResObservation) for o in bl.observations):
        raise NotImplementedError("LowResObservation runs in float32 only")
    packed = pack_blends([bl], want_spectra=False, dtype=np.float64)
    if packed["sched"] is not None or packed["prior"] is not None:
        raise NotImplementedError("declared schedules and priors run in float32 only")
    C, Ny, Nx = fr.shape
    kernels, ker_on = [], []
    shapes = {o._diff_kernels.image.shape[1:] for o in bl.observations if o._diff_kernels is not None}
    if len(shapes) > 1:
        raise NotImplementedError("float64 loop: the difference kernels of all observations must share one shape")
    kshape = shapes.pop() if shapes else None
    for o in bl.observations:
        Co = o.images.shape[0]
        if o._diff_kernels is None:
            ker_on += [0] * Co
            if kshape:
                kernels += [np.zeros(kshape)] * Co
        else:
            ker_on += [1] * Co
            kernels += list(np.broadcast_to(o._diff_kernels.image, (Co,) + kshape))
    out = _native.fit_f64(C, Ny, Nx, packed["sed"], packed["morph"], packed["cflags"], packed["center"], packed["shift"],
                          packed["conv"], packed["images"], packed["weights"], packed["plane_band"], packed["n_obs"],
                          np.array(kernels) if kshape else None, ker_on, _centroid_weight(fr), packed["it0"][0],
                          packed["prev_loss"][0], packed["src_first"], max_iter, e_rel, approximate_L)
    n = out["n_iter"]
    res = dict(sed=out["sed"], morph=out["morph"], center=out["center"], shift=out["shift"], conv=out["conv"],
               n_iter=np.array([n]), logL=out["logL"][None] if n else None, lipschitz=out["lipschitz"][None])
    unpack_blends([bl], packed, res)
    return dict(steps=n, n_blends=1, shape=fr.shape, dtype="float64")


def fit_many(blends, max_iter=200, e_rel=1e-2, approximate_L=False, check_every=8, stream=None):
    """`Blend.fit` for a list of independent blends, batched on one GPU.
    Returns the list of timing dicts (one per frame-shape group)."""
    groups = {}
    infos = []
    for bl in blends:
        if bl.frame.dtype == np.float64:
            infos.append(fit_f64(bl, max_iter, e_rel, approximate_L))
            continue
        layout = tuple((o._band_slice.start, o._band_slice.stop, o.images.shape[0], getattr(o, "lr_shape", None))
                       for o in bl.observations)
        groups.setdefault((bl.frame.shape, layout), []).append(bl)
    for (shape, _), group in groups.items():
        packed = pack_blends(group)
        bt = _native.Batch(stream=stream, **packed)
        try:
            info = bt.fit(max_iter, e_rel, approximate_L, check_every)
            out = bt.download()
        finally:
            bt.close()
        unpack_blends(group, packed, out)
        info.update(n_blends=len(group), shape=shape)
        infos.append(info)
    return infos


class Blend(ComponentTree):
    """The blended scene (blend.py:11-223)."""

    def __init__(self, sources, observations):
        ComponentTree.__init__(self, sources)
        try:
            iter(observations)
        except TypeError:
            observations = (observations,)
        self.observations = observations
        self.mse = []

    @property
    def it(self):
        """Number of iterations run in the `fit` method"""
        return len(self.mse)

    @property
    def converged(self):
        for c in self.components:
            if (c.flags & (BlendFlag.SED_NOT_CONVERGED | BlendFlag.MORPH_NOT_CONVERGED)).value > 0:
                return False
        return True

    def fit(self, max_iter=200, e_rel=1e-2, approximate_L=False):
        """Fit the model for each source to the data (blend.py:65-102), on the GPU."""
        fit_many([self], max_iter, e_rel, approximate_L)
