"""Generate tests/golden/ref_golden.npz by running the reference package
(/root/reference, imported unmodified through ref_shims.py) in the build
container.  The reference cannot travel to the GPU box, so its outputs are
committed as fixtures.  Run:  python tests/golden/make_golden.py
"""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_shims  # noqa: E402

scarlet = ref_shims.install()
from scarlet.cache import Cache  # noqa: E402

out = {}
rng = np.random.default_rng(1234)

# ---- monotonic weights + sweeps (operator.py:540-621, operators_pybind11.cc:27-50)
for tag, shape, center in [("a", (9, 7), (2, 5)), ("b", (16, 13), (3, 3)), ("c", (12, 12), (6, 6))]:
    w = scarlet.operator.getRadialMonotonicWeights(shape, useNearest=False, center=center)
    out["mono_%s_shape" % tag] = np.array(shape)
    out["mono_%s_center" % tag] = np.array(center)
    out["mono_%s_weights" % tag] = w
    out["mono_%s_didx" % tag] = scarlet.operator.sort_by_radius(shape, center)
    for dt in (np.float32, np.float64):
        x = rng.random(shape).astype(dt)
        out["mono_%s_in_%s" % (tag, np.dtype(dt).name)] = x.copy()
        for th in (0, .1):
            y = x.copy()
            scarlet.operator.prox_strict_monotonic(shape, use_nearest=False, thresh=th, center=center)(y, 0)
            out["mono_%s_out_%s_t%d" % (tag, np.dtype(dt).name, int(th * 10))] = y

# ---- symmetry (operator.py:231-350)
for tag, shape, center, shift in [("a", (11, 14), (3, 8), (0.3, -0.2)), ("b", (21, 21), (12, 9), (-0.45, 0.05)),
                                  ("c", (16, 16), (5, 5), (0.0, 0.0)), ("d", (15, 15), (7, 7), (0.2, 0.1)),
                                  ("e", (40, 59), (30, 11), (0.123, 0.377))]:
    x = rng.random(shape)
    x[rng.random(shape) < .1] = 0
    out["sym_%s_in" % tag] = x.copy()
    out["sym_%s_center" % tag] = np.array(center)
    out["sym_%s_shift" % tag] = np.array(shift)
    y = x.copy()
    scarlet.operator.prox_uncentered_symmetry(y, 0, center, "kspace", None, shift)
    out["sym_%s_kspace" % tag] = y          # in-place result, as update.symmetric sees it
    y = x.copy()
    scarlet.operator.prox_uncentered_symmetry(y, 0, center, "soft", None, None, 1)
    out["sym_%s_soft" % tag] = y
    y = x.copy()
    scarlet.operator.prox_uncentered_symmetry(y, 0, center, "sdss")
    out["sym_%s_sdss" % tag] = y

# ---- centre measurement (measurement.py)
psf = scarlet.psf.generate_psf_image(scarlet.psf.gaussian, (41, 41), sigma=.9).image
out["psf_target_41"] = psf
for tag, shape, center in [("a", (30, 40), (3, 20)), ("b", (64, 64), (40, 22)), ("c", (25, 25), (12, 23))]:
    m = rng.random(shape).astype(np.float32)
    yy, xx = np.indices(shape)
    m += 5 * np.exp(-((yy - center[0] - .3) ** 2 + (xx - center[1] + .4) ** 2) / 8).astype(np.float32)
    out["meas_%s_in" % tag] = m
    out["meas_%s_center" % tag] = np.array(center)
    out["meas_%s_maxpix" % tag] = np.array(scarlet.measurement.max_pixel(m, center))
    c, s = scarlet.measurement.psf_weighted_centroid(m, psf.astype(np.float32), center)
    out["meas_%s_centroid" % tag] = np.array(c)
    out["meas_%s_shift" % tag] = np.array(s, dtype=np.float64)


# ---- reference tests/test_blend.py:7-52 scenario
def init_data(shape, coords, amplitudes, dtype):
    import scipy.signal
    B, Ny, Nx = shape
    K = len(coords)
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
    images = np.array([scipy.signal.convolve(img, p, method="direct", mode="same")
                       for img, p in zip(images, psfs)], dtype=dtype)
    psfs /= psfs.sum(axis=(1, 2))[:, None, None]
    return target_psf, psfs, images


shape = (6, 31, 55)
coords = [(20, 10), (10, 30), (17, 42)]
target_psf, psfs, images = init_data(shape, coords, [3, 2, 1], np.float64)
out["tb_target_psf"], out["tb_psfs"], out["tb_images"] = target_psf, psfs, images
frame = scarlet.Frame(images.shape, psfs=target_psf[None], dtype=np.float64)
obs = scarlet.Observation(images, psfs=psfs).match(frame)
out["tb_diff_kernels"] = obs._diff_kernels.image
srcs = [scarlet.PointSource(frame, c, obs) for c in coords]
blend = scarlet.Blend(srcs, obs)
out["tb_point_model"] = obs.render(blend.get_model())
blend.fit(10)
out["tb_point_mse"] = np.array(blend.mse)
Cache._cache.clear()
srcs = [scarlet.ExtendedSource(frame, c, obs, np.ones(6)) for c in coords]
blend = scarlet.Blend(srcs, obs)
out["tb_ext_init_seds"] = np.array([c.sed for c in blend.components])
out["tb_ext_init_morphs"] = np.array([c.morph for c in blend.components])
blend.fit(100)
out["tb_ext_mse"] = np.array(blend.mse)
out["tb_ext_seds"] = np.array([c.sed for c in blend.components])
out["tb_ext_morphs"] = np.array([c.morph for c in blend.components])
out["tb_ext_centers"] = np.array([s.pixel_center for s in srcs])


# ---- config 1: data/hsc_cosmos.npz, docs/quickstart.ipynb cells 3-14
def run_npz(tag, fn, bg, max_iter, star_mode=False):
    Cache._cache.clear()
    d = np.load(os.path.join("/root/reference/data", fn))
    images, psfs, cat = d["images"], d["psfs"], d["catalog"]
    mpsf = scarlet.psf.generate_psf_image(scarlet.psf.gaussian, psfs[0].shape, sigma=.9).image
    frame = scarlet.Frame(images.shape, psfs=mpsf[None])
    obs = scarlet.Observation(images, psfs=psfs).match(frame)
    bg_rms = np.ones(len(images), dtype=images.dtype) * bg
    srcs = []
    if star_mode:
        seen = set()
        for row in cat:
            key = (int(row["y"]), int(row["x"]))
            if key in seen:
                continue
            seen.add(key)
            if row["is_star"]:
                srcs.append(scarlet.PointSource(frame, key, obs, fix_morph=True))
            else:
                srcs.append(scarlet.ExtendedSource(frame, key, obs, bg_rms))
    else:
        srcs = [scarlet.ExtendedSource(frame, (row["y"], row["x"]), obs, bg_rms) for row in cat]
    blend = scarlet.Blend(srcs, obs)
    out[tag + "_diff_kernels"] = obs._diff_kernels.image
    out[tag + "_init_seds"] = np.array([c.sed for c in blend.components])
    out[tag + "_init_morphs"] = np.array([c.morph for c in blend.components])
    out[tag + "_init_centers"] = np.array([s.pixel_center for s in srcs])
    out[tag + "_init_shifts"] = np.array([s.shift for s in srcs], dtype=np.float64)
    blend.fit(max_iter)
    out[tag + "_mse"] = np.array(blend.mse, dtype=np.float64)
    out[tag + "_seds"] = np.array([c.sed for c in blend.components])
    out[tag + "_morphs"] = np.array([c.morph for c in blend.components])
    out[tag + "_centers"] = np.array([s.pixel_center for s in srcs])
    out[tag + "_model"] = blend.get_model()
    print(tag, "iterations", blend.it, "converged", blend.converged, "mse", blend.mse[0], blend.mse[-1])


run_npz("hsc", "hsc_cosmos.npz", 0.1, 200)
run_npz("sim", "psf_unmatched_sim.npz", 20, 200, star_mode=True)

np.savez_compressed(os.path.join(HERE, "ref_golden.npz"), **out)
print("wrote", len(out), "arrays")

# ---- input fixtures: the reference's data/*.npz cannot be read on the GPU box
for fn, outn, bg, star in [("hsc_cosmos.npz", "hsc_cosmos_inputs.npz", 0.1, False),
                           ("psf_unmatched_sim.npz", "psf_unmatched_sim_inputs.npz", 20, True)]:
    d = np.load("/root/reference/data/" + fn)
    coords, stars, seen = [], [], set()
    for row in d["catalog"]:
        key = (int(row["y"]), int(row["x"]))
        if star:
            if key in seen:
                continue
            seen.add(key)
            stars.append(bool(row["is_star"]))
        else:
            stars.append(False)
        coords.append(key)
    np.savez_compressed(os.path.join(HERE, outn), images=d["images"], psfs=d["psfs"],
                        coords=np.array(coords), is_star=np.array(stars), bg_rms=bg)
