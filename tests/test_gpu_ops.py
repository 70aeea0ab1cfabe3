"""GPU parity of the array-level operators, called through the C ABI, against
the oracle and the reference fixtures."""
import numpy as np
import pytest
from numpy.testing import assert_array_equal, assert_allclose

from oracle import ref_prox, ref_fft, ref_blend

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nat():
    from scarlet_b200 import _native
    _native.require_device()
    return _native


def test_monotonic_kat(nat):
    """reference tests/test_operator.py:30-56 through the CUDA wavefront (f64 and f32)."""
    X = np.arange(25).reshape(1, 5, 5).astype(np.float64)
    Y = X.copy()
    nat.prox_weighted_monotonic(Y, [(2, 2)], 0.25)
    gold = X[0].copy()
    ref_prox.build_monotonic((5, 5), None, .25)(gold)
    assert_array_equal(Y[0], gold)
    Y = X.copy()
    nat.prox_monotonic_nearest(Y, [(2, 2)], 0.0)
    assert_array_equal(Y[0], [[0, 1, 2, 3, 4], [5, 6, 7, 8, 9], [10, 11, 12, 12, 12],
                              [11, 12, 12, 12, 12], [12, 12, 12, 12, 12]])


@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_monotonic_vs_reference_fixture(nat, golden, tag):
    shape = tuple(golden["mono_%s_shape" % tag])
    center = tuple(golden["mono_%s_center" % tag])
    for dt in ("float32", "float64"):
        for th, t in ((0, 0), (.1, 1)):
            x = golden["mono_%s_in_%s" % (tag, dt)].copy()[None]
            nat.prox_weighted_monotonic(x, [center], th)
            assert_array_equal(x[0], golden["mono_%s_out_%s_t%d" % (tag, dt, t)])   # bit-exact


# (90, 80) has streamed sweep plans with 13 pixel bits; (100, 100) and (128, 87) exceed the plan entry format and
# take the generic level loop; (2, 3) is the smallest shape the sweep accepts
@pytest.mark.parametrize("shape", [(64, 64), (69, 57), (40, 59), (7, 100), (90, 80), (100, 100), (128, 87), (2, 3)])
def test_monotonic_random_centres_bit_exact(nat, shape):
    rng = np.random.default_rng(3)
    n = 24 if shape[0] * shape[1] < 8000 else 8
    centers = np.stack([rng.integers(0, shape[0], n), rng.integers(0, shape[1], n)], axis=1)
    centers[0] = (0, 0)
    centers[1] = (shape[0] - 1, shape[1] - 1)
    centers[2] = (shape[0] // 2, shape[1] // 2)
    for dt in (np.float32, np.float64):
        x = (rng.random((n,) + shape) - .2).astype(dt)
        y = x.copy()
        nat.prox_weighted_monotonic(y, centers, 0.1 if dt == np.float64 else 0.0)
        for i in range(n):
            g = x[i].copy()
            ref_prox.build_monotonic(shape, tuple(centers[i]), 0.1 if dt == np.float64 else 0.0)(g)
            assert_array_equal(y[i], g)
    # float32 with the threshold of source initialisation (source.py:165-166)
    x = (rng.random((n,) + shape) - .2).astype(np.float32)
    y = x.copy()
    nat.prox_weighted_monotonic(y, centers, 0.1)
    for i in range(n):
        g = x[i].copy()
        ref_prox.build_monotonic(shape, tuple(centers[i]), 0.1)(g)
        assert_array_equal(y[i], g)


@pytest.mark.parametrize("tag", ["a", "b", "c", "d", "e"])
def test_symmetry_vs_reference_fixture(nat, golden, tag):
    x64 = golden["sym_%s_in" % tag]
    c = tuple(golden["sym_%s_center" % tag])
    shift = tuple(golden["sym_%s_shift" % tag])
    x = x64.astype(np.float32)[None]
    for mode, key in (("soft", "soft"), ("sdss", "sdss")):
        y = x.copy()
        nat.prox_symmetry(y, [c], None, mode, 1.0)
        g = x[0].copy()
        ref_prox.prox_uncentered_symmetry(g, 0, c, mode, None, None, 1)
        assert_array_equal(y[0], g)                                     # f32 oracle: bit-exact
        assert_allclose(y[0], golden["sym_%s_%s" % (tag, key)], rtol=2e-7)   # the reference's f64 run
    y = x.copy()
    nat.prox_symmetry(y, [c], [shift], "kspace")
    gold = x64.astype(np.float32).astype(np.float64)
    sl = ref_prox.uncentered_slices(gold.shape, c)
    if sl is None:
        gold = ref_prox.prox_kspace_symmetry(gold, 0, shift)     # what a caller of the operator gets
    else:
        gold[sl] = ref_prox.prox_kspace_symmetry(gold[sl], 0, shift)
    assert_allclose(y[0], gold, rtol=0, atol=2e-6 * np.abs(gold).max())


def test_symmetry_random(nat):
    rng = np.random.default_rng(5)
    shape, n = (64, 64), 16
    x = rng.random((n,) + shape).astype(np.float32)
    x[rng.random(x.shape) < .2] = 0
    centers = np.stack([rng.integers(3, 61, n), rng.integers(3, 61, n)], axis=1)
    shifts = rng.uniform(-.5, .5, (n, 2))
    y = x.copy()
    nat.prox_symmetry(y, centers, shifts, "kspace")
    for i in range(n):
        g = x[i].astype(np.float64)
        ref_prox.prox_uncentered_symmetry(g, 0, tuple(centers[i]), "kspace", None, tuple(shifts[i]))
        assert_allclose(y[i], g, rtol=0, atol=3e-6)


@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_measure_vs_reference_fixture(nat, golden, tag):
    m = golden["meas_%s_in" % tag][None]
    c = tuple(golden["meas_%s_center" % tag])
    cc, _ = nat.measure(m, [c], None, True, False)
    assert_array_equal(cc[0], golden["meas_%s_maxpix" % tag])
    cc, ss = nat.measure(m, [c], golden["psf_target_41"].astype(np.float32), False, True)
    assert_array_equal(cc[0], golden["meas_%s_centroid" % tag])
    assert_allclose(ss[0], golden["meas_%s_shift" % tag], rtol=0, atol=2e-6)


def test_convolve_vs_oracle(nat, golden):
    """Observation.render on the reference's test_blend scene (tests/test_blend.py:56-73), f32."""
    rng = np.random.default_rng(7)
    dk = golden["tb_diff_kernels"].astype(np.float32)
    C, Ny, Nx = 6, 31, 55
    model = rng.random((2, C, Ny, Nx)).astype(np.float32)
    spec = nat.kernel_spectrum(dk, Ny, Nx)
    idx = np.tile(np.arange(C), 2)
    out = nat.convolve(model, spec, idx)
    adj = nat.convolve(model, spec, idx, adjoint=True)
    for i in range(2):
        ref = ref_fft.convolve(model[i].astype(np.float64), dk.astype(np.float64), axes=(1, 2))
        assert_allclose(out[i], ref, rtol=0, atol=1e-5 * np.abs(ref).max())
        ref = ref_fft.correlate(model[i].astype(np.float64), dk.astype(np.float64), axes=(1, 2))
        assert_allclose(adj[i], ref, rtol=0, atol=1e-5 * np.abs(ref).max())
    out = nat.convolve(model, None, -np.ones(2 * C))
    assert_array_equal(out, model)


def test_convolve_shapes(nat):
    rng = np.random.default_rng(8)
    for (Ny, Nx, k) in [(64, 64, 41), (69, 57, 41), (40, 59, 31), (3, 5, 7), (64, 64, 16)]:
        ker = rng.random((1, k, k))
        model = rng.random((1, 1, Ny, Nx)).astype(np.float32)
        out = nat.convolve(model, nat.kernel_spectrum(ker, Ny, Nx), [0])
        ref = ref_fft.convolve(model[0].astype(np.float64), ker, axes=(1, 2))
        assert_allclose(out[0], ref, rtol=0, atol=1e-5 * np.abs(ref).max())


# ---------------------------------------------------------------- images beyond the shared-memory limit
def test_large_image_operators(nat):
    """The array-level operators on a 300x280 image (config 4 class): the monotonic level loop, the symmetry
    operators and the measurements work on the image in HBM; the convolution runs as overlap-save tiles."""
    rng = np.random.default_rng(9)
    shape, n = (300, 280), 3
    centers = np.array([(152, 139), (40, 250), (299, 0)])      # none at the exact frame centre (operator.py:203-204)
    for dt, th in ((np.float32, 0.0), (np.float32, 0.1), (np.float64, 0.1)):
        x = (rng.random((n,) + shape) - .2).astype(dt)
        y = x.copy()
        nat.prox_weighted_monotonic(y, centers, th)
        for i in range(n):
            g = x[i].copy()
            ref_prox.build_monotonic(shape, tuple(centers[i]), th)(g)
            assert_array_equal(y[i], g)                                  # bit-exact
    yy, xx = np.indices(shape)
    x = np.array([np.exp(-np.hypot(yy - cy - .3, xx - cx + .2) / 6.) for cy, cx in centers[:2]]).astype(np.float32)
    x += (.01 * rng.random(x.shape)).astype(np.float32)
    shifts = np.array([(.3, -.2), (-.4, .1)])
    for mode in ("soft", "sdss"):
        y = x.copy()
        nat.prox_symmetry(y, centers[:2], None, mode, 1.0)
        for i in range(2):
            g = x[i].copy()
            ref_prox.prox_uncentered_symmetry(g, 0, tuple(centers[i]), mode, None, None, 1)
            assert_array_equal(y[i], g)
    y = x.copy()
    nat.prox_symmetry(y, centers[:2], shifts, "kspace")
    for i in range(2):
        g = x[i].astype(np.float64)
        ref_prox.prox_uncentered_symmetry(g, 0, tuple(centers[i]), "kspace", None, tuple(shifts[i]))
        assert_allclose(y[i], g, rtol=0, atol=1e-5)
    psf = np.exp(-np.hypot(*np.indices((41, 41)) - 20.) ** 2 / 2).astype(np.float32)
    cc, ss = nat.measure(x, centers[:2] + 1, psf, True, True)
    for i in range(2):
        c0 = ref_prox.max_pixel(x[i], tuple(centers[i] + 1))
        c1, s1 = ref_prox.psf_weighted_centroid(x[i], psf, c0)
        assert_array_equal(cc[i], c1)
        assert_allclose(ss[i], s1, rtol=0, atol=5e-6)
    ker = rng.random((2, 41, 41))
    model = rng.random((1, 2, 300, 280)).astype(np.float32)
    spec = nat.kernel_spectrum(ker, 300, 280)
    h = nat.kernel_half(41, 41, 300, 280)
    out = nat.convolve(model, spec, [0, 1], ker_half=h)
    adj = nat.convolve(model, spec, [0, 1], adjoint=True, ker_half=h)
    ref = ref_fft.convolve(model[0].astype(np.float64), ker, axes=(1, 2))
    assert_allclose(out[0], ref, rtol=0, atol=1e-5 * np.abs(ref).max())
    ref = ref_fft.correlate(model[0].astype(np.float64), ker, axes=(1, 2))
    assert_allclose(adj[0], ref, rtol=0, atol=1e-5 * np.abs(ref).max())


def _moffat(shape, alpha, beta, dy=0.0, dx=0.0):
    y, x = np.indices(shape, dtype=np.float64)
    r2 = (y - shape[0] // 2 - dy) ** 2 + (x - shape[1] // 2 - dx) ** 2
    p = (1 + r2 / alpha ** 2) ** -beta
    return p / p.sum()


@pytest.mark.parametrize("tag", ["tb", "hsc", "sim"])
def test_match_psfs_vs_reference_fixture(nat, golden, tag):
    """Difference kernels of the reference's own runs (Observation.match in tests/test_blend.py and the two
    notebooks' data sets, tests/golden/make_golden.py) through scl_match_psfs_f32."""
    import os
    from scarlet_b200 import psf as spsf
    gdir = os.path.join(os.path.dirname(__file__), "golden")
    if tag == "tb":
        psfs, target = golden["tb_psfs"], golden["tb_target_psf"]
    else:
        fn = {"hsc": "hsc_cosmos_inputs.npz", "sim": "psf_unmatched_sim_inputs.npz"}[tag]
        psfs = np.load(os.path.join(gdir, fn))["psfs"]
        target = spsf.generate_psf_image(spsf.gaussian, psfs[0].shape, sigma=.9).image
    want = golden[tag + "_diff_kernels"]
    diff, spec = nat.match_psfs(psfs[None], target[None], frame_shape=(60, 51))
    scale = np.abs(want).max()
    assert_allclose(diff[0], want, rtol=0, atol=1e-5 * scale)
    # the spectrum is that of the float32 kernel the same call returned, placed like the host helper does
    host = nat.kernel_spectrum(diff[0], 60, 51)
    assert_allclose(spec[0], host, rtol=0, atol=2e-6 * np.abs(host).max())


@pytest.mark.parametrize("ky,kx,kym,kxm", [(41, 41, 41, 41), (25, 27, 15, 15), (24, 24, 41, 41), (31, 30, 11, 12)])
def test_match_psfs_batch_vs_oracle(nat, ky, kx, kym, kxm):
    """Many PSF sets in one launch (odd, even and mixed stamp sizes; one model PSF per set or one shared) against the
    float64 oracle (oracle/ref_fft.match_psfs, fft.py:282-301)."""
    rng = np.random.default_rng(ky * 100 + kxm)
    n, C = 9, 5
    obs = np.stack([[_moffat((ky, kx), rng.uniform(2.0, 3.5), rng.uniform(2.5, 4.0), *rng.uniform(-.3, .3, 2))
                     for _ in range(C)] for _ in range(n)]).astype(np.float32)
    shared = _moffat((kym, kxm), 1.3, 4.5)[None].astype(np.float32)
    each = np.stack([_moffat((kym, kxm), rng.uniform(1.1, 1.5), 4.5) for _ in range(n)]).astype(np.float32)
    for model in (shared, each):
        diff, spec = nat.match_psfs(obs, model, frame_shape=(48, 64))
        for i in range(n):
            m = model[0 if len(model) == 1 else i]
            want = ref_fft.match_psfs(obs[i].astype(np.float64), m[None].astype(np.float64))
            assert_allclose(diff[i], want, rtol=0, atol=1e-5 * np.abs(want).max())
        host = nat.kernel_spectrum(diff.reshape(n * C, ky, kx), 48, 64).reshape(spec.shape)
        assert_allclose(spec, host, rtol=0, atol=2e-6 * np.abs(host).max())
    only_diff, none = nat.match_psfs(obs, shared)
    assert none is None
    assert_array_equal(only_diff, nat.match_psfs(obs, shared, frame_shape=(48, 64))[0])
    with pytest.raises(ValueError):
        nat.match_psfs(obs, each[:3])


def test_bbox_kats():
    """reference tests/test_bbox.py:98-151 (trim, flux_at_edge) through scl_trim_f32."""
    from scarlet_b200 import bbox
    x = np.arange(25).reshape(5, 5)
    x[0] = 0
    x[:, -2:] = 0
    assert bbox.trim(x) == bbox.Box((1, 0), 4, 3)
    x += 10
    assert bbox.trim(x) == bbox.Box((0, 0), 5, 5)
    assert bbox.trim(x, 10) == bbox.Box((1, 0), 4, 3)
    with pytest.raises(ValueError):
        bbox.trim(x, 100)

    x = np.arange(25).reshape(5, 5)
    assert bbox.flux_at_edge(x) is True
    for sl, want in [((0,), True), ((-1,), True), ((slice(None), 0), True), ((slice(None), -1), False)]:
        x[sl] = 0
        assert bbox.flux_at_edge(x) is want
    x = np.arange(25).reshape(5, 5)
    assert bbox.flux_at_edge(x, 5) is True
    for sl, want in [((0,), True), ((-1,), True), ((slice(None), 0), True), ((slice(None), -1), False)]:
        x[sl] = 5
        assert bbox.flux_at_edge(x, 5) is want


def test_trim_many_vs_numpy(nat):
    """Bounds, edge flags and pixel sums of many ragged footprints at once against numpy (bbox.py:170-210)."""
    rng = np.random.default_rng(3)
    n, Ny, Nx = 40, 37, 52
    X = np.zeros((n, Ny, Nx), np.float32)
    for i in range(n - 2):
        y0, x0 = rng.integers(0, Ny - 3), rng.integers(0, Nx - 3)
        y1, x1 = rng.integers(y0 + 1, Ny + 1), rng.integers(x0 + 1, Nx + 1)
        X[i, y0:y1, x0:x1] = rng.random((y1 - y0, x1 - x0)) * (rng.random((y1 - y0, x1 - x0)) > .6)
    X[n - 2] = -1.0                                      # nothing above the cut
    for mv in (0.0, 0.35):
        bounds, edge, tot = nat.trim(X, mv)
        for i in range(n):
            nz = np.where(X[i] > mv)
            if len(nz[0]) == 0:
                assert_array_equal(bounds[i], [-1] * 4)
            else:
                assert_array_equal(bounds[i], [nz[0].min(), nz[0].max(), nz[1].min(), nz[1].max()])
            e = max(X[i][:, 0].max(), X[i][:, -1].max(), X[i][0].max(), X[i][-1].max()) > mv
            assert bool(edge[i]) == bool(e)
        assert_allclose(tot, X.astype(np.float64).sum(axis=(1, 2)), rtol=1e-12, atol=1e-9)


class _Comp:
    """The attributes of a Component the update functions touch."""
    def __init__(self, sed, morph, L_morph=1.0):
        self.sed, self.morph, self.L_morph = sed, morph, L_morph
        self.shape = (len(sed),) + morph.shape

    @property
    def step_morph(self):
        return 1 / self.L_morph


def test_sparsity_kats():
    """reference tests/test_update.py:74-97 (sparse_l0, sparse_l1) through scl_prox_sparse_f32."""
    from scarlet_b200 import update
    sed = np.arange(6)
    morph = np.arange(25, dtype=float).reshape(5, 5)
    src = _Comp(sed.copy(), morph.copy(), L_morph=1)
    update.sparse_l0(src, thresh=4)
    want = morph.copy()
    want[0, :-1] = 0
    assert_array_equal(src.sed, sed)
    assert_array_equal(src.morph, want)
    src = _Comp(sed.copy(), morph.astype(np.float32), L_morph=0.5)
    update.sparse_l1(src, thresh=2)
    want = np.zeros(25)
    want[5:] = np.arange(20) + 1
    assert_array_equal(src.morph, want.reshape(5, 5))
    # both signs
    x = np.array([[-3., -1., 0., .5, 2.]], dtype=np.float32)
    a = _Comp(sed, x.copy())
    update.sparse_l1(a, 1.0)
    assert_array_equal(a.morph, [[-2., -0., 0., 0., 1.]])
    a = _Comp(sed, x.copy())
    update.sparse_l0(a, 1.0)
    assert_array_equal(a.morph, [[-3., -1., 0., 0., 2.]])


def _np_threshold(morph):
    """measurement.py:96-113 restated with numpy (np.int -> int)."""
    m = morph[morph > 0]
    bins = 50
    if m.size < 500:
        bins = max(int(m.size / 10), 1)
        if bins == 1:
            return 0, bins
    hist, edges = np.histogram(np.log10(m).reshape(-1), bins)
    cut = np.where(hist == 0)[0]
    if len(cut) == 0:
        return 0, bins
    return 10 ** edges[cut[-1]], bins


def test_threshold_kat():
    """reference tests/test_update.py:99-119: a 7x7 Gaussian blob on a uniform noise floor keeps only the blob."""
    from scarlet_b200 import update, measurement, psf as spsf, bbox
    rs = np.random.RandomState(0)
    noise = rs.rand(21, 21) * 2
    signal = np.zeros(noise.shape)
    signal[7:14, 7:14] = spsf.generate_psf_image(spsf.gaussian, (21, 21), normalize=False, amplitude=10,
                                                 sigma=3)[7:14, 7:14].image
    morph = (signal + noise).astype(np.float32)
    sed = np.arange(5)
    src = _Comp(sed.copy(), morph.copy())
    thresh, nbins = measurement.threshold(src.morph)
    want_t, want_b = _np_threshold(morph)
    assert nbins == want_b == 44
    assert_allclose(thresh, want_t, rtol=1e-5)
    want = np.zeros(morph.shape)
    want[7:14, 7:14] = morph[7:14, 7:14]
    update.threshold(src)
    assert_array_equal(src.sed, sed)
    assert_array_equal(src.morph, want)
    assert src.bboxes["thresh"] == bbox.Box((7, 7), 7, 7)


def test_threshold_many_vs_numpy(nat):
    """Cuts, bin counts, zeroed images and boxes of a stack of morphologies against numpy.histogram: bright cores on
    faint noise floors (a gap in the histogram), plain noise (no gap), tiny footprints (fewer bins, or none)."""
    rng = np.random.default_rng(8)
    n, Ny, Nx = 24, 40, 33
    X = np.zeros((n, Ny, Nx), np.float32)
    y, x = np.indices((Ny, Nx))
    for i in range(n):
        if i % 4 == 0:
            X[i] = rng.random((Ny, Nx))                                  # no structure
        elif i % 4 == 1:
            X[i, 5:8, 4:4 + 1 + i // 4] = rng.random((3, 1 + i // 4)) + .1      # 3 .. 18 pixels
        else:
            core = 50 * np.exp(-((y - 17 - i % 5) ** 2 + (x - 15) ** 2) / 8.0)
            core[core < 5] = 0
            X[i] = core + 1e-3 * rng.random((Ny, Nx)) * (rng.random((Ny, Nx)) > .3)
    th, bins = nat.threshold(X.copy())
    Y = X.copy()
    th2, bins2, bounds = nat.threshold(Y, apply=True)
    assert_array_equal(th2, th)
    assert_array_equal(bins2, bins)
    gaps = 0
    for i in range(n):
        wt, wb = _np_threshold(X[i])
        assert bins[i] == wb
        assert_allclose(th[i], wt, rtol=1e-5, atol=0)
        gaps += wt > 0
        keep = X[i].astype(np.float64) >= th[i]
        assert_array_equal(Y[i], np.where(keep, X[i], 0))
        nz = np.where(Y[i] > 0)
        assert_array_equal(bounds[i], [nz[0].min(), nz[0].max(), nz[1].min(), nz[1].max()])
    assert gaps >= n // 2


@pytest.fixture(scope="module")
def gold_resample():
    import os
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_golden_resample.npz"))


@pytest.mark.parametrize("name", ["lanczos", "bilinear", "catmull_rom", "mitchel_netravali", "cubic_spline",
                                  "quintic_spline"])
def test_fft_resample_vs_reference_fixture(gold_resample, name):
    """interpolation.fft_resample of the reference itself (tests/golden/make_golden_resample.py), every kernel of
    interpolation.py:139-270, six sub-pixel shifts; taps equal to 1e-14, images to float32 rounding."""
    from scarlet_b200 import interpolation
    g = gold_resample
    fn = getattr(interpolation, name)
    for taps, d in zip(g["taps_" + name], (.217, -.691, .5, 0.)):
        y, w = fn(d)
        assert_allclose(np.concatenate([y, w]), taps, rtol=0, atol=1e-14)
    img = g["img"]
    for (dy, dx), want in zip(g["shifts"], g["resample_" + name]):
        got = interpolation.fft_resample(img, dy, dx, kernel=fn)
        assert got.shape == img.shape and got.dtype == img.dtype
        assert_allclose(got, want, rtol=0, atol=3e-7 * np.abs(want).max())
    with pytest.raises(ValueError):
        interpolation.lanczos(1.5)


def test_fft_resample_kat():
    """reference tests/test_interpolation.py:365-391 (bilinear shifts of a 6x6 ramp)."""
    from scarlet_b200 import interpolation
    small = np.arange(36).reshape(6, 6)
    img = np.zeros((11, 11))
    img[2:8, 2:8] = small
    for (dy, dx), (oy, ox) in [((.217, -.026), (3, 1)), ((-.691, .321), (1, 3))]:
        Dy, Dx = abs(dy), abs(dx)
        truth = np.zeros(img.shape)
        truth[2:8, 2:8] += small * (1 - Dx) * (1 - Dy)
        truth[2:8, ox:ox + 6] += small * Dx * (1 - Dy)
        truth[oy:oy + 6, 2:8] += small * (1 - Dx) * Dy
        truth[oy:oy + 6, ox:ox + 6] += small * Dx * Dy
        assert_allclose(interpolation.fft_resample(img, dy, dx, kernel=interpolation.bilinear), truth, rtol=0, atol=1e-5)


def test_translation_and_threshold_vs_reference_fixture(gold_resample):
    """update.translation (both directions) and update.threshold / measurement.threshold as run by the reference."""
    from scarlet_b200 import update, measurement
    g = gold_resample
    for direction, want in zip((1, -1), g["translation"]):
        c = _Comp(np.arange(3), g["img"].astype(np.float32))
        c.shift = (.31, -.44)
        update.translation(c, direction)
        assert_allclose(c.morph, want, rtol=0, atol=3e-7 * np.abs(want).max())
    c = _Comp(np.arange(3), g["thresh_in"].copy())
    t, b = measurement.threshold(c.morph)
    assert b == int(g["thresh_bins"])
    assert_allclose(t, float(g["thresh_value"]), rtol=1e-5)
    update.threshold(c)
    assert_array_equal(c.morph, g["thresh_out"])
    bb = c.bboxes["thresh"]
    assert_array_equal([bb.bottom, bb.top, bb.left, bb.right], g["thresh_box"])


def test_errors(nat):
    x = np.zeros((1, 5, 5), np.float32)
    with pytest.raises(ValueError):
        nat.prox_weighted_monotonic(x, [(7, 0)])
    with pytest.raises(TypeError):
        nat.prox_weighted_monotonic(x.astype(np.int32), [(2, 2)])
    with pytest.raises(NotImplementedError):
        nat.kernel_spectrum(np.zeros((1, 125, 125)), 64, 64)      # kernel wider than a convolution tile
    assert nat.kernel_spectrum(np.zeros((1, 41, 41)), 512, 512).shape == (1, 128, 65)     # big frames: tiled
    y = np.zeros((0, 5, 5), np.float32)
    nat.prox_weighted_monotonic(y, np.zeros((0, 2)))          # empty batch is a no-op
