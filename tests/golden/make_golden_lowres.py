"""Generate tests/golden/ref_golden_lowres.npz: the reference's multi-resolution path (LowResObservation,
observation.py:242-599; resampling.match_patches; interpolation.sinc_interp / mk_shifter; CombinedExtendedSource,
source.py:183-240,495-535) run UNMODIFIED from /root/reference through ref_shims.py, on a synthetic
3-band 40x40 high-resolution + 5-band 12x12 low-resolution scene (pixel scale 4, unrotated, sub-pixel grid offset)
described by the affine WCS stand-in wcs_stub.py (astropy is absent here).  Run in the build container:
  python tests/golden/make_golden_lowres.py

Extra compat patch (besides ref_shims): numpy >= 1.24 refuses `np.array((a, b))` for arrays of different length
(observation.py:374 wraps the two 1-D shifters like that); the stand-in returns the object array old numpy made.
Case R exercises the reference's rotated branch (10 degrees): it needs every low-resolution pixel inside the
high-resolution frame (observation.py:486-489 reshape) and a cold operator cache (mk_shifter's cache key ignores
`real`, interpolation.py:319; repro_rotated_lowres.py shows both outcomes).

`autograd.grad` stand-in for the low-resolution observation: the adjoint of the reference's own linear `_render`,
taken from its dense matrix (one `_render` call per model pixel).
"""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_shims  # noqa: E402
from wcs_stub import make_pair  # noqa: E402

scarlet = ref_shims.install()
import autograd  # noqa: E402
import autograd.numpy as agnp  # noqa: E402
from scarlet.cache import Cache  # noqa: E402

_arr = np.array


def ragged_array(obj, *a, **k):
    try:
        return _arr(obj, *a, **k)
    except ValueError:
        out = np.empty(len(obj), dtype=object)
        for i, o in enumerate(obj):
            out[i] = o
        return out


agnp.array = ragged_array
psf = scarlet.psf
CH = ["h0", "h1", "h2", "l0", "l1", "l2", "l3", "l4"]


def scene(hr_shape, lr_shape, seed, offset=(0.3, -0.2), angle=0.0):
    """frame, hr observation, lr observation, true model, true centres"""
    Cache._cache.clear()
    rng = np.random.default_rng(seed)
    (HY, HX), (LY, LX) = hr_shape, lr_shape
    whr, wlr = make_pair(hr_shape, lr_shape, scale=4.0, angle_deg=angle, lr_offset=offset)
    mpsf = psf.generate_psf_image(psf.gaussian, (15, 15), sigma=.9).image.astype(np.float32)
    frame = scarlet.Frame((8, HY, HX), wcs=whr, psfs=mpsf[None], channels=CH)
    hr_psfs = np.array([psf.generate_psf_image(psf.gaussian, (15, 15), sigma=1.2 + .15 * b).image
                        for b in range(3)]).astype(np.float32)
    lr_psfs = np.array([psf.generate_psf_image(psf.gaussian, (9, 9), sigma=.8 + .05 * b).image
                        for b in range(5)]).astype(np.float32)
    centers = [(HY * .35, HX * .4), (HY * .6, HX * .62), (HY * .3, HX * .72)]
    yy, xx = np.indices((HY, HX))
    model = np.zeros((8, HY, HX), np.float32)
    for (cy, cx) in centers:
        re, q, pa = rng.uniform(1.5, 3.5), rng.uniform(.6, 1), rng.uniform(0, np.pi)
        u = (xx - cx) * np.cos(pa) + (yy - cy) * np.sin(pa)
        v = (-(xx - cx) * np.sin(pa) + (yy - cy) * np.cos(pa)) / q
        morph = np.exp(-np.hypot(u, v) / re)
        sed = rng.dirichlet(np.ones(8)) * 10 ** rng.uniform(2.5, 3.5)
        model += (sed[:, None, None] * morph[None]).astype(np.float32)
    # images: the reference's own forward models of the true scene + noise
    ohr = scarlet.Observation(np.zeros((3, HY, HX), np.float32), psfs=hr_psfs, wcs=whr, channels=CH[:3]).match(frame)
    olr = scarlet.LowResObservation(np.zeros((5, LY, LX), np.float32), wcs=wlr, psfs=lr_psfs, channels=CH[3:])
    olr.match(frame)
    hr_img = (ohr.render(model) + rng.normal(0, 1, (3, HY, HX))).astype(np.float32)
    lr_img = rng.normal(0, 1, (5, LY, LX)).astype(np.float32)
    y0, y1 = int(olr._coord_lr[0].min()), int(olr._coord_lr[0].max())
    x0, x1 = int(olr._coord_lr[1].min()), int(olr._coord_lr[1].max())
    lr_img[:, y0:y1 + 1, x0:x1 + 1] += olr._render(model)
    w_lr = rng.uniform(.5, 1.5, (5, y1 - y0 + 1, x1 - x0 + 1)).astype(np.float32)
    Cache._cache.clear()
    ohr = scarlet.Observation(hr_img, psfs=hr_psfs, wcs=whr, channels=CH[:3]).match(frame)
    olr = scarlet.LowResObservation(lr_img, wcs=wlr, psfs=lr_psfs, weights=w_lr, channels=CH[3:])
    olr.match(frame)
    return frame, ohr, olr, model, centers, dict(hr_psfs=hr_psfs, lr_psfs=lr_psfs, mpsf=mpsf, hr_img=hr_img,
                                                 lr_img=lr_img, w_lr=w_lr, offset=np.array(offset),
                                                 angle=np.float64(angle))


def dump_match(tag, olr, out):
    out[tag + "fft_shape"] = np.array(olr._fft_shape)
    out[tag + "lr_shape"] = np.array(olr.lr_shape)
    out[tag + "small_axis"] = np.array(bool(olr.small_axis))
    out[tag + "coord_lr0"], out[tag + "coord_lr1"] = np.array(olr._coord_lr[0]), np.array(olr._coord_lr[1])
    out[tag + "coord_hr0"], out[tag + "coord_hr1"] = np.array(olr._coord_hr[0]), np.array(olr._coord_hr[1])
    out[tag + "resconv_op"] = olr._resconv_op.astype(np.float32)
    out[tag + "other_shifts"] = np.array(olr.other_shifts, dtype=np.float64)


def dense_operator(olr, frame):
    """[C_lr, n_lr, Ny*Nx]: the reference's `_render` applied to every unit model (it is linear)."""
    C, Ny, Nx = frame.shape
    Cl = olr.frame.C
    n = int(np.prod(olr.lr_shape))
    L = np.zeros((Cl, n, Ny * Nx), np.float64)
    e = np.zeros((C, Ny, Nx), frame.dtype)
    for p in range(Ny * Nx):
        e[olr._band_slice].reshape(Cl, -1)[:, p] = 1
        L[:, :, p] = olr._render(e).reshape(Cl, n)
        e[olr._band_slice].reshape(Cl, -1)[:, p] = 0
    return L


out = {}
# ---- case A: square frames (small_axis True), with a fit
frame, ohr, olr, truth, centers, inp = scene((40, 40), (12, 12), seed=11)
for k, v in inp.items():
    out["A_" + k] = v
dump_match("A_", olr, out)
out["A_diff_psf"] = olr.build_diffkernel(frame, None).image.astype(np.float32)
rng = np.random.default_rng(3)
model = (truth * rng.uniform(.8, 1.2, truth.shape)).astype(np.float32)
out["A_model"] = model
out["A_render"] = olr._render(model)
out["A_loss"] = np.float64(olr.get_loss(model))
L = dense_operator(olr, frame)
y0, x0 = int(olr._coord_lr[0].min()), int(olr._coord_lr[1].min())
ny, nx = olr.lr_shape
sub = olr.images[:, y0:y0 + ny, x0:x0 + nx]
r = (olr.weights ** 2 * (olr._render(model) - sub)).reshape(5, -1).astype(np.float64)
out["A_grad"] = np.einsum("cq,cqp->cp", r, L).reshape(5, 40, 40)

# autograd stand-in for the two-observation blend
_grad_hr = autograd.grad


def grad(fun, argnums):
    blend = fun.__self__

    def g(*params):
        import scarlet.fft as rfft
        K = blend.K
        seds, morphs = params[:K], params[K:]
        m = blend.get_model(seds, morphs)
        total = 0
        G = np.zeros(m.shape, dtype=np.float64)
        for obs in blend.observations:
            total = total + obs.get_loss(m)
            if obs is olr:
                rr = (obs.weights ** 2 * (obs._render(m) - sub)).reshape(5, -1).astype(np.float64)
                G[obs._band_slice] += np.einsum("cq,cqp->cp", rr, L).reshape(5, 40, 40)
            else:
                rr = obs.weights * obs.weights * (obs.render(m) - obs.images)
                rr = rfft._kspace_operation(rfft.Fourier(rr), obs._diff_kernels, 3,
                                            lambda a, b: a * np.conj(b), rr.shape, axes=(1, 2)).image
                G[obs._band_slice] += rr
        G = G.astype(m.dtype)
        blend.mse.append(total)
        sg = [np.sum(G * morphs[k][None], axis=(1, 2)) for k in range(K)]
        mg = [np.sum(seds[k][:, None, None] * G, axis=0) for k in range(K)]
        return tuple(sg + mg)
    return g


autograd.grad = grad
import scarlet.blend as sblend  # noqa: E402
sblend.grad = grad
bg = [np.ones(3, np.float32), np.ones(5, np.float32)]
srcs = []
for (cy, cx) in centers:
    sky = (round(cy) - 40 / 2, round(cx) - 40 / 2)          # wcs_stub: sky = (world_y, world_x)
    srcs.append(scarlet.CombinedExtendedSource(frame, sky, [ohr, olr], bg, obs_idx=0))
blend = scarlet.Blend(srcs, [ohr, olr])
out["A_init_seds"] = np.array([c.sed for c in blend.components])
out["A_init_morphs"] = np.array([c.morph for c in blend.components])
out["A_init_centers"] = np.array([s.pixel_center for s in srcs])
N_IT = 12
blend.fit(N_IT, e_rel=1e-9)
out["A_mse"] = np.array(blend.mse, dtype=np.float64)
out["A_seds"] = np.array([c.sed for c in blend.components])
out["A_morphs"] = np.array([c.morph for c in blend.components])
out["A_centers"] = np.array([s.pixel_center for s in srcs])
out["A_fit_model"] = blend.get_model()
print("A: fft", olr._fft_shape, "lr", olr.lr_shape, "small_axis", olr.small_axis, "loss", out["A_loss"],
      "it", blend.it, "mse", blend.mse[0], blend.mse[-1])

# ---- case B: the low-resolution frame inside the high-resolution one, other offset; match + render + loss only
# (unrotated frames must be square: resampling.py:90-92 multiplies per-column and per-row masks, so
#  `small_axis` (observation.py:470) is always True on the path that works)
frame, ohr, olr, truth, centers, inp = scene((48, 48), (10, 10), seed=12, offset=(-0.15, 0.4))
for k, v in inp.items():
    out["B_" + k] = v
dump_match("B_", olr, out)
model = (truth * rng.uniform(.8, 1.2, truth.shape)).astype(np.float32)
out["B_model"] = model
out["B_render"] = olr._render(model)
out["B_loss"] = np.float64(olr.get_loss(model))
print("B: fft", olr._fft_shape, "lr", olr.lr_shape, "small_axis", olr.small_axis, "loss", out["B_loss"])

# ---- case R: the low-resolution frame rotated by 10 degrees and entirely inside the high-resolution one (the
# reference's rotated branch, observation.py:443-454,483-503; it runs with a cold operator cache, see
# repro_rotated_lowres.py): match + render + loss + adjoint
frame, ohr, olr, truth, centers, inp = scene((64, 64), (10, 10), seed=13, offset=(0.25, -0.1), angle=10.0)
assert olr.isrot
for k, v in inp.items():
    out["R_" + k] = v
dump_match("R_", olr, out)
out["R_diff_psf"] = olr.build_diffkernel(frame, (olr.cos_rot, olr.sin_rot)).image.astype(np.float32)
model = (truth * rng.uniform(.8, 1.2, truth.shape)).astype(np.float32)
out["R_model"] = model
out["R_render"] = olr._render(model)
out["R_loss"] = np.float64(olr.get_loss(model))
L = dense_operator(olr, frame)
y0, x0 = int(olr._coord_lr[0].min()), int(olr._coord_lr[1].min())
ny, nx = olr.lr_shape
sub = olr.images[:, y0:y0 + ny, x0:x0 + nx]
r = (olr.weights ** 2 * (olr._render(model) - sub)).reshape(5, -1).astype(np.float64)
out["R_grad"] = np.einsum("cq,cqp->cp", r, L).reshape(5, 64, 64)
print("R: fft", olr._fft_shape, "lr", olr.lr_shape, "small_axis", olr.small_axis, "loss", out["R_loss"])
np.savez_compressed(os.path.join(HERE, "ref_golden_lowres.npz"), **out)
print(os.path.getsize(os.path.join(HERE, "ref_golden_lowres.npz")) / 1e6, "MB")
