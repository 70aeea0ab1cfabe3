"""Shared test helpers: pack an oracle Blend into the arrays of scarlet_b200._native.Batch."""
import os
import numpy as np

from oracle import ref_blend as rb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def npz_blend(tag, dtype=np.float32):
    """Oracle scene for configs 1 ("hsc") and 2 ("sim"), built from the input fixtures."""
    fn = {"hsc": "hsc_cosmos_inputs.npz", "sim": "psf_unmatched_sim_inputs.npz"}[tag]
    d = np.load(os.path.join(GOLD, fn))
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


def multi_blend(dtype=np.float32):
    """config 1 with its first rows as MultiComponentSource (fixture ref_golden_multi.npz), oracle objects."""
    d = np.load(os.path.join(GOLD, "hsc_cosmos_inputs.npz"))
    g = np.load(os.path.join(GOLD, "ref_golden_multi.npz"))
    images, psfs = d["images"], d["psfs"]
    mpsf = rb.generate_psf_image(rb.gaussian, psfs[0].shape, sigma=.9)
    frame = rb.Frame(images.shape, psfs=mpsf[None], dtype=dtype)
    obs = rb.Observation(images, psfs=psfs).match(frame)
    bg = np.ones(len(images), dtype=images.dtype) * float(d["bg_rms"])
    srcs = [(rb.MultiComponentSource if i < int(g["n_multi"]) else rb.ExtendedSource)(frame, (y, x), obs, bg)
            for i, (y, x) in enumerate(d["coords"])]
    return g, frame, obs, srcs, rb.Blend(srcs, obs)


def synthetic_blend(seed, C=5, Ny=64, Nx=64, kmin=2, kmax=8, psf_size=41, dtype=np.float32, _cache={}):
    """One synthetic blend of BASELINE config 3, oracle objects."""
    rng = np.random.default_rng(20260411 + seed)
    key = (C, psf_size)
    if key not in _cache:
        mpsf = rb.generate_psf_image(rb.gaussian, (psf_size, psf_size), sigma=.9)
        psfs = np.array([rb.generate_psf_image(rb.gaussian, (psf_size, psf_size), sigma=1.2 + .15 * b)
                         for b in range(C)])
        _cache[key] = (mpsf, psfs)
    mpsf, psfs = _cache[key]
    K = int(rng.integers(kmin, kmax + 1))
    centers = []
    while len(centers) < K:
        c = rng.uniform(12, min(Ny, Nx) - 12, 2)
        if all(np.hypot(*(c - o)) >= 6 for o in centers):
            centers.append(c)
    yy, xx = np.indices((Ny, Nx))
    images = np.zeros((C, Ny, Nx))
    for c in centers:
        re, n = rng.uniform(1.5, 5), rng.uniform(.7, 2.5)
        q, pa = rng.uniform(.5, 1), rng.uniform(0, np.pi)
        dy, dx = yy - c[0], xx - c[1]
        u = dx * np.cos(pa) + dy * np.sin(pa)
        v = (-dx * np.sin(pa) + dy * np.cos(pa)) / q
        morph = np.exp(-(np.hypot(u, v) / re) ** (1 / n))
        sed = rng.dirichlet(np.ones(C)) * 10 ** rng.uniform(np.log10(50), np.log10(5000))
        from scipy.signal import fftconvolve
        for b in range(C):
            images[b] += sed[b] * fftconvolve(morph, psfs[b], mode="same")
    images += rng.normal(0, 1, images.shape)
    images = images.astype(dtype)
    weights = np.ones(images.shape, dtype=dtype)
    frame = rb.Frame(images.shape, psfs=mpsf[None], dtype=dtype)
    obs = rb.Observation(images, psfs=psfs.astype(dtype), weights=weights).match(frame)
    bg = np.ones(C, dtype=dtype)
    srcs = []
    for c in centers:
        try:
            srcs.append(rb.ExtendedSource(frame, (int(round(c[0])), int(round(c[1]))), obs, bg))
            srcs[-1].pixel_center_init = (int(round(c[0])), int(round(c[1])))
        except rb.SourceInitError:
            pass
    return frame, obs, srcs, rb.Blend(srcs, obs)


def multi_obs_blend(seed, split=((0, 3), (1, 5)), **kw):
    """A synthetic blend seen by several band-sliced observations (Blend._loss sums them, blend.py:135-139;
    Observation._band_slice, observation.py:184-189): observation i holds model channels split[i][0]..split[i][1]-1,
    the later ones with wider PSFs and their own noise.  Sources are initialised on the full-band observation."""
    frame, obs, srcs, _ = synthetic_blend(seed, **kw)
    C, Ny, Nx = frame.shape
    chan = tuple("b%d" % c for c in range(C))
    frame.channels = chan
    rng = np.random.default_rng(777 + seed)
    ps = obs.frame.psfs.shape[-1]
    many = []
    for i, (lo, hi) in enumerate(split):
        psfs = np.array([rb.generate_psf_image(rb.gaussian, (ps, ps), sigma=1.2 + .15 * b + .2 * i) for b in range(lo, hi)])
        # the full-band images re-blurred is not needed for a parity test: reuse the bands, add fresh noise
        img = (obs.images[lo:hi] + rng.normal(0, .5 * i, (hi - lo, Ny, Nx))).astype(frame.dtype)
        w = None if i % 2 else np.full(img.shape, 1. / (1 + i), dtype=frame.dtype)
        many.append(rb.Observation(img, psfs=psfs.astype(frame.dtype), weights=w, channels=chan[lo:hi]).match(frame))
    return frame, many, srcs, rb.Blend(srcs, many)


def lowres_scene(gold, tag, dtype=np.float32):
    """Oracle objects of the multi-resolution fixture (tests/golden/make_golden_lowres.py): model frame,
    high-resolution Observation, LowResObservation."""
    sys_path_golden()
    from wcs_stub import make_pair
    from oracle import ref_lowres as rl
    hr_img, lr_img = gold[tag + "hr_img"], gold[tag + "lr_img"]
    angle = float(gold[tag + "angle"]) if tag + "angle" in gold else 0.0
    whr, wlr = make_pair(hr_img.shape[1:], lr_img.shape[1:], scale=4.0, angle_deg=angle,
                         lr_offset=tuple(gold[tag + "offset"]))
    chan = ["h0", "h1", "h2", "l0", "l1", "l2", "l3", "l4"]
    frame = rb.Frame((8,) + hr_img.shape[1:], psfs=gold[tag + "mpsf"][None], channels=chan, dtype=dtype)
    frame.wcs = whr
    ohr = rb.Observation(hr_img, psfs=gold[tag + "hr_psfs"], channels=chan[:3])
    ohr.frame.wcs = whr
    ohr.match(frame)
    olr = rl.LowResObservation(lr_img, wcs=wlr, psfs=gold[tag + "lr_psfs"], weights=gold[tag + "w_lr"],
                               channels=chan[3:]).match(frame)
    return frame, ohr, olr


def lowres_blend(gold, frame, ohr, olr):
    """CombinedExtendedSource per true centre of fixture A + the two-observation Blend (oracle objects)."""
    from oracle import ref_lowres as rl
    HY, HX = frame.shape[1:]
    bg = [np.ones(3, np.float32), np.ones(5, np.float32)]
    srcs = []
    for (cy, cx) in [(HY * .35, HX * .4), (HY * .6, HX * .62), (HY * .3, HX * .72)]:
        sky = (round(cy) - HY / 2, round(cx) - HX / 2)
        srcs.append(rl.CombinedExtendedSource(frame, sky, [ohr, olr], bg, obs_idx=0))
    return srcs, rb.Blend(srcs, [ohr, olr])


def scene_blend(size, sources, bands=6, psf_size=41, seed=None, multi_fraction=.7):
    """Oracle objects for one config-4-class scene of scarlet_b200.synth.scene_arrays (70 % two-component sources
    sharing a centre like MultiComponentSource, the rest ExtendedSource, every constraint on), started from the
    generator's own profiles.  Returns (batch arrays as generated, extra, oracle Blend)."""
    from scarlet_b200 import synth
    kw = {} if seed is None else dict(seed=seed)
    batch, extra = synth.scene_arrays(size, sources, bands, psf_size, multi_fraction=multi_fraction, **kw)
    C, Ny, Nx = batch["C"], batch["Ny"], batch["Nx"]
    frame = rb.Frame((C, Ny, Nx), psfs=extra["mpsf"], dtype=np.float32)
    obs = rb.Observation(extra["images"], psfs=extra["psfs"], weights=batch["weights"][0]).match(frame)
    cw = rb._centroid_weight(frame)
    srcs, k, K = [], 0, len(batch["sed"])
    while k < K:
        if extra["multi"][k]:
            m = object.__new__(rb.MultiComponentSource)
            m.frame, m.symmetric, m.monotonic, m.parent = frame, True, True, None
            m.pixel_center, m.shift, m.centroid_weight = tuple(batch["center"][k]), (0., 0.), cw
            m.components = [rb.Component(frame, batch["sed"][j], batch["morph"][j]) for j in (k, k + 1)]
            srcs.append(m)
            k += 2
        else:
            s_ = object.__new__(rb.ExtendedSource)
            rb.Component.__init__(s_, frame, batch["sed"][k], batch["morph"][k])
            s_.symmetric, s_.monotonic, s_.centroid_weight = True, True, cw
            s_.pixel_center, s_.shift = tuple(batch["center"][k]), (0., 0.)
            srcs.append(s_)
            k += 1
    return batch, extra, rb.Blend(srcs, obs)


def seed_oracle(blend, out, k0, logL):
    """Put an oracle Blend into the state a device run left behind: components [k0, k0 + K) of a downloaded batch
    (`out` of _native.Batch.download) and the loss history `logL` (one entry per iteration run).  The previous
    iterate the next convergence check compares with is the state itself (blend.py:180-182)."""
    k = k0
    for s in blend.sources:
        for c in s.components:
            c.sed = out["sed"][k].astype(blend.frame.dtype).copy()
            c.morph = out["morph"][k].astype(blend.frame.dtype).copy()
            c.flags = int(out["conv"][k])
            c.last_sed, c.last_morph = c.sed.copy(), c.morph.copy()
            k += 1
        if hasattr(s, "pixel_center"):
            s.pixel_center = (int(out["center"][k - 1, 0]), int(out["center"][k - 1, 1]))
            s.shift = (float(out["shift"][k - 1, 0]), float(out["shift"][k - 1, 1]))
    blend.mse = [float(v) for v in logL]
    return k


def sys_path_golden():
    import sys
    if GOLD not in sys.path:
        sys.path.insert(0, GOLD)


def pack(blends):
    """Arrays for _native.Batch from oracle Blend objects sharing one frame shape."""
    from scarlet_b200 import _native as nat
    C, Ny, Nx = blends[0].frame.shape
    comp_off, sed, morph, cflags, center, shift, conv, src_first = [0], [], [], [], [], [], [], []
    images, weights, specs, ker_idx, cws, cw_idx, it0, prev = [], [], [], [], [], [], [], []
    spec_cache, cw_cache = {}, {}
    any_w = any(isinstance(o.weights, np.ndarray) for b in blends for o in b.observations)
    is_lr = lambda o: hasattr(o, "dense_operator")       # oracle.ref_lowres.LowResObservation
    plane_band = [c for o in blends[0].observations if not is_lr(o) for c in range(C)[o.band_slice]]
    plane_band += [c for o in blends[0].observations if is_lr(o) for c in range(C)[o.band_slice]]
    lr_images, lr_weights, lr_ops, lr_op_idx, lr_cache = [], [], [], [], {}
    ker_half = max([nat.kernel_half(*o.diff_kernels.shape[1:], Ny, Nx) for b in blends for o in b.observations
                    if not is_lr(o) and o.diff_kernels is not None] + [0])
    for bl in blends:
        for s in bl.sources:
            first = len(sed)
            for c in s.components:
                f = 0
                if isinstance(s, rb.PointSource):
                    f |= nat.F_CONSTRAINED | nat.F_KSPACE
                    if s.symmetric:
                        f |= nat.F_SYMMETRIC
                    if s.monotonic:
                        f |= nat.F_MONOTONIC
                elif isinstance(s, rb.MultiComponentSource):
                    f |= nat.F_CONSTRAINED | nat.F_SHARED_CENTER
                    if s.symmetric:
                        f |= nat.F_SYMMETRIC
                    if s.monotonic:
                        f |= nat.F_MONOTONIC
                src_first.append(first)
                if c.fix_sed:
                    f |= nat.F_FIX_SED
                if c.fix_morph:
                    f |= nat.F_FIX_MORPH
                cflags.append(f)
                sed.append(c.sed)
                morph.append(c.morph)
                center.append(getattr(s, "pixel_center", (Ny // 2, Nx // 2)))
                shift.append(getattr(s, "shift", (0., 0.)))
                conv.append(c.flags)
        comp_off.append(len(sed))
        for obs in bl.observations:
            if not is_lr(obs):
                continue
            ys, xs = obs._window()
            sub = obs.images[:, ys, xs]
            lr_images.append(sub.reshape(len(sub), -1))
            w = obs.weights if isinstance(obs.weights, np.ndarray) else np.ones_like(sub)
            lr_weights.append(np.broadcast_to(w, sub.shape).reshape(len(sub), -1))
            if id(obs) not in lr_cache:
                lr_cache[id(obs)] = len(lr_ops)
                lr_ops.append(obs.dense_operator(bl.frame.shape).astype(np.float32))
            lr_op_idx.append(lr_cache[id(obs)])
        for obs in bl.observations:
            if is_lr(obs):
                continue
            Co = len(obs.images)
            images.extend(obs.images)
            if any_w:
                weights.extend(obs.weights if isinstance(obs.weights, np.ndarray) else np.ones_like(obs.images))
            if obs.diff_kernels is None:
                ker_idx += [-1] * Co
            else:
                key = obs.diff_kernels.tobytes()
                if key not in spec_cache:
                    spec_cache[key] = sum(len(sp) for sp in specs)
                    specs.append(nat.kernel_spectrum(obs.diff_kernels, Ny, Nx))
                ker_idx += [spec_cache[key] + c for c in range(Co)]
        cw = rb._centroid_weight(bl.frame).astype(np.float32)
        key = cw.tobytes()
        if key not in cw_cache:
            cw_cache[key] = len(cws)
            cws.append(cw)
        cw_idx.append(cw_cache[key])
        it0.append(bl.it)
        prev.append(bl.mse[-1] if bl.mse else 0.)
    return dict(C=C, Ny=Ny, Nx=Nx, comp_off=np.array(comp_off), sed=np.array(sed), morph=np.array(morph),
                cflags=np.array(cflags), center=np.array(center), shift=np.array(shift, dtype=np.float64),
                conv=np.array(conv), images=np.array(images), weights=np.array(weights) if any_w else None,
                kernel_spec=np.concatenate(specs) if specs else None, ker_idx=np.array(ker_idx),
                cw=np.array(cws), cw_idx=np.array(cw_idx), it0=np.array(it0), prev_loss=np.array(prev),
                src_first=np.array(src_first), ker_half=ker_half,
                **(dict(plane_band=np.array(plane_band, dtype=np.int32), n_obs=len(blends[0].observations))
                   if len(blends[0].observations) > 1 or plane_band != list(range(C)) else {}),
                **(dict(lr=dict(images=np.array(lr_images), weights=np.array(lr_weights), op=np.array(lr_ops),
                                op_idx=np.array(lr_op_idx))) if lr_images else {}))
