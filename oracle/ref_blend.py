"""Oracle: scene model, likelihood, analytic gradient, Lipschitz step,
constraint schedule and convergence of `Blend.fit` (TEST INFRASTRUCTURE ONLY).

Follows /root/reference/scarlet:
  Frame / Observation.match / render / get_loss   observation.py:29-54,122-239
  Component / ComponentTree.get_model             component.py:91-112,148-170,321-347
  Blend.fit / _backward / _loss / _check_convergence / _set_lipschitz   blend.py:65-223
  PointSource / ExtendedSource / MultiComponentSource (+ update schedules)
                                                  source.py:139-180,242-291,334-492,551-641
  generate_psf_image / gaussian                   psf.py:21-89, interpolation.py:497-552

The reference differentiates `_loss` with the third-party `autograd` package
(blend.py:44-45), which is absent here.  Its result is the exact adjoint of the
forward operator; `Observation.grad` below restates it analytically (SURVEY
A.3) and tests/test_oracle_blend.py checks it against torch.autograd in float64
and against finite differences.
"""
import numpy as np
from . import ref_fft, ref_prox

SED_NOT_CONVERGED = 1
MORPH_NOT_CONVERGED = 2


# ---------------------------------------------------------------- psf.py
def gaussian(y, x, y0=0, x0=0, amplitude=None, sigma=1):
    if amplitude is None:
        amplitude = 1 / (np.pi ** 2 * sigma ** 2)
    X, Y = np.meshgrid(x, y)
    return amplitude * np.exp(-((X - x0) ** 2 + (Y - y0) ** 2) / (2 * sigma ** 2))


def generate_psf_image(func, shape, subsamples=10, normalize=True, **kwargs):
    """psf.py:55-89 with interpolation.py:497-552 (2-D trapezoid rule, including
    the reference's 0.4 factor)."""
    ry, rx = np.array(shape) // 2
    y = np.linspace(-ry, ry, shape[0])
    x = np.linspace(-rx, rx, shape[1])
    dy, dx = y[1] - y[0], x[1] - x[0]
    n = subsamples
    _x = np.linspace(x[0] - dx / 2, x[-1] + dx / 2, len(x) * n + 1)
    _y = np.linspace(y[0] - dy / 2, y[-1] + dy / 2, len(y) * n + 1)
    z = func(_y, _x, **kwargs)
    dz = 0.4 * (z[:-1, :-1] + z[1:, :-1] + z[:-1, 1:] + z[1:, 1:])
    vol = dy * dx * dz / n / n
    _ny, _nx = len(_y) // n, len(_x) // n
    vol = np.array(np.split(np.array(np.split(vol, _nx, axis=1)), _ny, axis=1)).sum(axis=(2, 3))
    if normalize:
        vol /= vol.sum()
    return vol


# ---------------------------------------------------------------- observation.py
class Frame:
    def __init__(self, shape, psfs=None, channels=None, dtype=np.float32):
        assert len(shape) == 3
        self.shape = tuple(shape)
        if psfs is not None:
            psfs = np.array(psfs)
            assert len(psfs) == 1 or len(psfs) == shape[0]
            if not np.allclose(psfs.sum(axis=(1, 2)), 1):
                psfs = psfs * (1 / psfs.sum(axis=(1, 2)))[:, None, None]
            if psfs.dtype != dtype:
                psfs = psfs.astype(dtype)
        self.psfs = psfs
        self.channels = channels
        self.dtype = dtype


class Observation:
    def __init__(self, images, psfs=None, weights=None, channels=None):
        self.frame = Frame(images.shape, psfs=psfs, channels=channels, dtype=images.dtype)
        self.images = np.array(images)
        self.weights = np.array(weights) if weights is not None else 1
        self.diff_kernels = None
        self.band_slice = slice(None)

    def match(self, model_frame):
        if self.frame.dtype != model_frame.dtype:
            self.frame.dtype = model_frame.dtype
            self.images = self.images.astype(model_frame.dtype)
            if isinstance(self.weights, np.ndarray):
                self.weights = self.weights.astype(model_frame.dtype)
            if self.frame.psfs is not None:
                self.frame.psfs = self.frame.psfs.astype(model_frame.dtype)
        self.band_slice = slice(None)
        if self.frame.channels is not model_frame.channels:
            bmin = list(model_frame.channels).index(self.frame.channels[0])
            bmax = list(model_frame.channels).index(self.frame.channels[-1])
            self.band_slice = slice(bmin, bmax + 1)
        self.diff_kernels = None
        if self.frame.psfs is not model_frame.psfs:
            self.diff_kernels = ref_fft.match_psfs(self.frame.psfs, model_frame.psfs)
        return self

    def render(self, model):
        m = model[self.band_slice]
        if self.diff_kernels is not None:
            m = ref_fft.convolve(m, self.diff_kernels, axes=(1, 2))
        return m

    def get_loss(self, model):
        return 0.5 * np.sum((self.weights * (self.render(model) - self.images)) ** 2)

    def grad(self, model):
        """d loss / d model (full model band axis; zero outside band_slice)."""
        r = self.weights * self.weights * (self.render(model) - self.images)
        if self.diff_kernels is not None:
            r = ref_fft.correlate(r, self.diff_kernels, axes=(1, 2))
        g = np.zeros(model.shape, dtype=model.dtype)
        g[self.band_slice] = r
        return g


# ---------------------------------------------------------------- component.py / source.py
class Component:
    symmetric = False
    monotonic = False

    def __init__(self, frame, sed, morph, fix_sed=False, fix_morph=False):
        self.frame = frame
        self.sed = np.array(sed, dtype=frame.dtype)
        self.morph = np.array(morph, dtype=frame.dtype)
        self.fix_sed, self.fix_morph = fix_sed, fix_morph
        self.flags = SED_NOT_CONVERGED | MORPH_NOT_CONVERGED
        self.last_sed = np.zeros(self.sed.shape, self.sed.dtype)
        self.last_morph = np.zeros(self.morph.shape, self.morph.dtype)
        self.parent = None

    @property
    def components(self):
        return [self]

    def get_model(self):
        return self.sed[:, None, None] * self.morph[None]

    def update(self):
        return self


class Prior:
    """component.py:39-67: gradient and Lipschitz callbacks evaluated on the component's parameters."""

    def __init__(self, grad_func, L_func):
        self._grad_func, self._L_func = grad_func, L_func
        self.sed_grad = self.morph_grad = 0

    def compute_grad(self, component):
        self.sed_grad, self.morph_grad = self._grad_func(component.sed, component.morph)
        self.L_sed, self.L_morph = self._L_func(component.sed, component.morph)


def _it(node):
    return 0 if node.parent is None else node.parent.it


def _symmetric(c, center, shift):
    """update.symmetric (update.py:170-199): the prox result is discarded, so the
    k-space prox is a no-op at the exact frame centre."""
    ref_prox.prox_uncentered_symmetry(c.morph, 0, center, "kspace", None, shift, .5)


def _monotonic(c, center):
    """update.monotonic (update.py:106-156), default thresh=0."""
    m = c.morph.copy()
    ref_prox.build_monotonic(m.shape, center, 0)(m)
    c.morph[:] = m


def _positive_normalized(c):
    ref_prox.prox_plus(c.sed)
    ref_prox.prox_plus(c.morph)
    ref_prox.normalized(c.sed, c.morph, "morph_max")


def _centroid_weight(frame):
    if frame.psfs is None:
        psf = generate_psf_image(gaussian, (41, 41), normalize=False, amplitude=1, sigma=.9)
        return psf / psf.max()
    return frame.psfs[0]


_MONO_CACHE = {}


def _monotonic_cached(shape, center, thresh=0):
    key = (tuple(shape), (int(center[0]), int(center[1])), thresh)
    if key not in _MONO_CACHE:
        _MONO_CACHE[key] = ref_prox.build_monotonic(shape, center, thresh)
    return _MONO_CACHE[key]


def _monotonic(c, center):  # noqa: F811  (cached variant, update.py:131-147)
    m = c.morph.copy()
    _monotonic_cached(m.shape, center, 0)(m)
    c.morph[:] = m


def get_psf_sed(coord, obs, frame):
    """source.py:41-71"""
    sed = obs.images[:, int(coord[0]), int(coord[1])].copy()
    if obs.frame.psfs is not None:
        sed /= obs.frame.psfs.max(axis=(1, 2))
    if frame.psfs is not None:
        sed = sed * frame.psfs[0].max()
    return sed


def build_detection_coadd(sed, bg_rms, obs, thresh=1):
    """source.py:101-136"""
    C = len(sed)
    if np.any(bg_rms <= 0):
        raise ValueError("bg_rms must be greater than zero in all channels")
    pos = [c for c in range(C) if sed[c] > 0]
    imgs = [obs.images[c] for c in pos]
    rms = np.array([bg_rms[c] for c in pos])
    w = np.array([sed[c] / bg_rms[c] ** 2 for c in pos])
    jac = np.array([sed[c] ** 2 / bg_rms[c] ** 2 for c in pos]).sum()
    detect = np.einsum('i,i...', w, imgs) / jac
    cutoff = thresh * np.sqrt((w ** 2 * rms ** 2).sum()) / jac
    return detect, cutoff


class SourceInitError(Exception):
    pass


def init_extended_source(coord, frame, obs, bg_rms, thresh=1., symmetric=True, monotonic=True):
    """source.py:139-180"""
    sed = get_psf_sed(coord, obs, frame)
    morph, cutoff = build_detection_coadd(sed, bg_rms, obs, thresh)
    center = (int(coord[0]), int(coord[1]))
    if symmetric:
        morph = ref_prox.prox_uncentered_symmetry(morph, 0, center=center, algorithm="sdss")
    if monotonic:
        morph = _monotonic_cached(morph.shape, center, .1)(morph).reshape(morph.shape)
    mask = morph > cutoff
    if mask.sum() == 0:
        raise SourceInitError("No flux above threshold={2} for source at y={0} x={1}".format(*coord, cutoff))
    morph[~mask] = 0
    morph /= morph[center]
    return sed, morph


def get_best_fit_seds(morphs, frame, obs):
    """source.py:74-98"""
    K = len(morphs)
    _m = morphs.reshape(K, -1)
    data = obs.images.reshape(obs.frame.shape[0], -1)
    return np.dot(np.linalg.inv(np.dot(_m, _m.T)), np.dot(_m, data.T))


def init_multicomponent_source(coord, frame, obs, bg_rms, flux_percentiles=None, thresh=1.,
                               symmetric=True, monotonic=True):
    """source.py:242-291"""
    if flux_percentiles is None:
        flux_percentiles = [25]
    sed, morph = init_extended_source(coord, frame, obs, bg_rms, thresh, symmetric, monotonic)
    K = len(flux_percentiles) + 1
    morphs = np.zeros((K,) + morph.shape, dtype=morph.dtype)
    morphs[0] = morph
    max_flux = morph.max()
    percs = np.sort(flux_percentiles)
    last = 0
    for k in range(1, K):
        ft = percs[k - 1] * max_flux / 100
        m = morph > ft
        morphs[k - 1][m] = ft - last
        morphs[k][m] = morph[m] - ft
        last = ft
    for k in range(K):
        morphs[k] /= morphs[k].max()
    return get_best_fit_seds(morphs, frame, obs), morphs


class PointSource(Component):
    """source.py:334-440"""

    def __init__(self, frame, coord, obs, symmetric=True, monotonic=True, **kw):
        C, Ny, Nx = frame.shape
        morph = np.zeros((Ny, Nx), obs.images.dtype)
        pixel = (int(coord[0]), int(coord[1]))
        if frame.psfs is None:
            morph[pixel] = 1
        else:
            psf = frame.psfs[0]
            sy, sx = (np.array(psf.shape) - 1) // 2
            cy, cx = (np.array(morph.shape) - 1) // 2
            bottom = int(pixel[0] - cy - sy) + (Ny >> 1)
            left = int(pixel[1] - cx - sx) + (Nx >> 1)
            top, right = bottom + psf.shape[0], left + psf.shape[1]
            ys = slice(max(0, bottom), min(Ny, top))
            iys = slice(max(0, -bottom), max(Ny - bottom, -top))
            xs = slice(max(0, left), min(Nx, right))
            ixs = slice(max(0, -left), max(Nx - left, -right))
            morph[ys, xs] = psf[iys, ixs]
        self.pixel_center = pixel
        sed = obs.images[:, pixel[0], pixel[1]].copy()
        if obs.frame.psfs is not None:
            sed /= obs.frame.psfs.max(axis=(1, 2))
        super().__init__(frame, sed, morph, **kw)
        self.symmetric, self.monotonic = symmetric, monotonic
        if symmetric:
            self.centroid_weight = _centroid_weight(frame)
        self.update()

    def update(self):
        it = _it(self)
        self.pixel_center = ref_prox.max_pixel(self.morph, self.pixel_center)
        if self.symmetric:
            if it % 5 == 0:
                self.pixel_center, self.shift = ref_prox.psf_weighted_centroid(
                    self.morph, self.centroid_weight, self.pixel_center)
            _symmetric(self, self.pixel_center, self.shift)
        if self.monotonic:
            _monotonic(self, self.pixel_center)
        _positive_normalized(self)
        return self


class ExtendedSource(PointSource):
    """source.py:443-492 (init always symmetric)"""

    def __init__(self, frame, coord, obs, bg_rms, thresh=1, symmetric=True, monotonic=True, **kw):
        self.symmetric, self.monotonic = symmetric, monotonic
        self.pixel_center = (int(coord[0]), int(coord[1]))
        sed, morph = init_extended_source(coord, frame, obs, bg_rms, thresh, True, monotonic)
        Component.__init__(self, frame, sed, morph, **kw)
        if symmetric:
            self.centroid_weight = _centroid_weight(frame)
        self.update()


class MultiComponentSource:
    """source.py:537-641"""

    def __init__(self, frame, coord, obs, bg_rms, thresh=1, flux_percentiles=None,
                 symmetric=True, monotonic=True):
        self.frame = frame
        self.symmetric, self.monotonic = symmetric, monotonic
        self.pixel_center = (int(coord[0]), int(coord[1]))
        seds, morphs = init_multicomponent_source(coord, frame, obs, bg_rms, flux_percentiles,
                                                  thresh, symmetric, monotonic)
        self.components = [Component(frame, seds[k], morphs[k]) for k in range(len(seds))]
        self.parent = None
        if symmetric:
            self.centroid_weight = _centroid_weight(frame)
        self.update()

    def update(self):
        it = _it(self)
        _morph = np.sum([c.morph * c.sed.sum() for c in self.components], axis=0)
        self.pixel_center = ref_prox.max_pixel(_morph, self.pixel_center)
        if self.symmetric and it % 5 == 0:
            self.pixel_center, self.shift = ref_prox.psf_weighted_centroid(
                _morph, self.centroid_weight, self.pixel_center)
        for c in self.components:
            if self.symmetric:
                # plain Components carry no `.shift` -> soft symmetry, strength 1
                ref_prox.prox_uncentered_symmetry(c.morph, 0, self.pixel_center, "kspace", None, None, .5)
            if self.monotonic:
                _monotonic(c, self.pixel_center)
            _positive_normalized(c)
        return self


# ---------------------------------------------------------------- blend.py
class Blend:
    """blend.py:11-223 with the analytic adjoint in place of autograd."""

    def __init__(self, sources, observations):
        if not hasattr(sources, "__iter__"):
            sources = (sources,)
        self.sources = tuple(sources)
        for s in self.sources:
            s.parent = self
        comps = []
        for s in self.sources:
            for c in s.components:
                if c not in comps:
                    comps.append(c)
        self.components = tuple(comps)
        try:
            iter(observations)
        except TypeError:
            observations = (observations,)
        self.observations = observations
        self.frame = self.components[0].frame
        self.mse = []

    @property
    def K(self):
        return len(self.components)

    @property
    def it(self):
        return len(self.mse)

    @property
    def converged(self):
        return all((c.flags & 3) == 0 for c in self.components)

    def get_model(self, seds=None, morphs=None):
        model = np.zeros(self.frame.shape, dtype=self.frame.dtype)
        for k, c in enumerate(self.components):
            if seds is not None:
                model = model + seds[k][:, None, None] * morphs[k][None]
            else:
                model = model + c.get_model()
        return model

    def backward(self):
        """blend.py:105-139: loss appended to mse, gradients of all seds/morphs."""
        model = self.get_model()
        total = 0
        G = np.zeros(model.shape, dtype=model.dtype)
        for obs in self.observations:
            total = total + obs.get_loss(model)
            G = G + obs.grad(model)
        self.mse.append(total)
        sg = [np.sum(G * c.morph[None], axis=(1, 2)) for c in self.components]
        mg = [np.sum(c.sed[:, None, None] * G, axis=0) for c in self.components]
        return sg, mg

    def set_lipschitz(self, approximate_L=False):
        """blend.py:186-223"""
        if approximate_L:
            LA = LS = 0
            for c in self.components:
                LS += (c.sed ** 2).sum().item()
                LA += (c.morph ** 2).sum().item()
            if self.it > 1 and self.mse[-1] > self.mse[-2]:
                LS *= 2
                LA *= 2
        else:
            seds = np.array([c.sed for c in self.components])
            S = np.array([c.morph for c in self.components]).reshape(self.K, -1)
            LA = np.real(np.linalg.eigvals(S.dot(S.T)).max())
            LS = np.real(np.linalg.eigvals(seds.T.dot(seds)).max())
        n = len(self.observations)
        return LA * n, LS * n      # (L_sed, L_morph)

    def check_convergence(self, e_rel):
        """blend.py:141-184"""
        e2 = e_rel ** 2
        if self.it > 1:
            conv = True
            for c in self.components:
                if ((c.last_sed - c.sed) ** 2).sum() <= e2 * (c.sed ** 2).sum():
                    c.flags &= ~SED_NOT_CONVERGED
                else:
                    c.flags |= SED_NOT_CONVERGED
                    conv = False
                if ((c.last_morph - c.morph) ** 2).sum() <= e2 * (c.morph ** 2).sum():
                    c.flags &= ~MORPH_NOT_CONVERGED
                else:
                    c.flags |= MORPH_NOT_CONVERGED
                    conv = False
        else:
            conv = False
        for c in self.components:
            c.last_sed = c.sed.copy()
            c.last_morph = c.morph.copy()
        return conv

    def step(self, e_rel=1e-2, approximate_L=False):
        """One pass of the loop body, blend.py:79-102.  Returns converged."""
        sg, mg = self.backward()
        self.L_sed, self.L_morph = self.set_lipschitz(approximate_L)
        for k, c in enumerate(self.components):
            c.L_sed, c.L_morph = self.L_sed, self.L_morph          # blend.py:88-89
            sgk, mgk = sg[k], mg[k]
            prior = getattr(c, "prior", None)
            if prior is not None:                                  # Component.backward_prior, component.py:177-187
                prior.compute_grad(c)
                if not c.fix_morph:
                    mgk = mgk + prior.morph_grad
                    c.L_morph = c.L_morph + prior.L_morph
                if not c.fix_sed:
                    sgk = sgk + prior.sed_grad
                    c.L_sed = c.L_sed + prior.L_sed
            if not c.fix_sed:
                c.sed = c.sed - (1 / c.L_sed) * sgk
            if not c.fix_morph:
                c.morph = c.morph - (1 / c.L_morph) * mgk
        for s in self.sources:
            s.update()
        return self.check_convergence(e_rel)

    def fit(self, max_iter=200, e_rel=1e-2, approximate_L=False):
        for _ in range(max_iter):
            if self.step(e_rel, approximate_L):
                break
