"""Sources: initialisation and per-iteration constraint schedule
(mirrors scarlet/source.py).

Initialisation is setup code that runs once per source; its array operators
(sdss symmetry is a flip/min on the host, monotonicity thresh=.1 and the first
constraint pass run on the device).  The per-iteration `update()` schedule
(source.py:402-440) is executed inside the CUDA loop by `Blend.fit`; calling
`update()` on a source object directly runs the same device kernel once.
"""
import logging

import numpy as np

from .component import Component, ComponentTree
from . import _native
from .psf import generate_psf_image, gaussian

logger = logging.getLogger("scarlet.source")


class SourceInitError(Exception):
    """Error during source initialization"""
    pass


def get_pixel_sed(sky_coord, observation):
    """SED at `sky_coord` in `observation` (source.py:21-38)."""
    pixel = observation.frame.get_pixel(sky_coord)
    return observation.images[:, pixel[0], pixel[1]].copy()


def get_psf_sed(sky_coord, observation, frame):
    """Pixel SED corrected for the per-band PSF peaks (source.py:41-71)."""
    sed = get_pixel_sed(sky_coord, observation)
    if observation.frame.psfs is not None:
        sed /= observation.frame.psfs.max(axis=(1, 2))
    if frame.psfs is not None:
        sed = sed * frame.psfs[0].max()
    return sed


def get_best_fit_seds(morphs, frame, observation):
    """Least-squares SEDs for fixed morphologies (source.py:74-98)."""
    K = len(morphs)
    _morph = morphs.reshape(K, -1)
    data = observation.images.reshape(observation.frame.C, -1)
    return np.dot(np.linalg.inv(np.dot(_morph, _morph.T)), np.dot(_morph, data.T))


def build_detection_coadd(sed, bg_rms, observation, thresh=1):
    """SED/variance weighted coadd and its noise cutoff (source.py:101-136)."""
    C = len(sed)
    if np.any(bg_rms <= 0):
        raise ValueError("bg_rms must be greater than zero in all channels")
    positive = [c for c in range(C) if sed[c] > 0]
    positive_img = [observation.images[c] for c in positive]
    positive_bgrms = np.array([bg_rms[c] for c in positive])
    weights = np.array([sed[c] / bg_rms[c] ** 2 for c in positive])
    jacobian = np.array([sed[c] ** 2 / bg_rms[c] ** 2 for c in positive]).sum()
    detect = np.einsum('i,i...', weights, positive_img) / jacobian
    bg_cutoff = thresh * np.sqrt((weights ** 2 * positive_bgrms ** 2).sum()) / jacobian
    return detect, bg_cutoff


def _sdss_symmetry(morph, center):
    """prox_uncentered_symmetry(..., algorithm="sdss") (operator.py:231-239,342-343): min(X, flip X) on the
    largest window centred on `center`.  Element-wise, exact in any precision; host numpy at init."""
    Ny, Nx = morph.shape
    py, px = center
    cy, cx = Ny // 2, Nx // 2
    if py == cy and px == cx:
        ys, xs = slice(None), slice(None)
    else:
        dy, dx = int(2 * (py - cy)), int(2 * (px - cx))
        if not Ny % 2:
            dy += 1
        if not Nx % 2:
            dx += 1
        xs = slice(None, dx) if dx < 0 else slice(dx, None)
        ys = slice(None, dy) if dy < 0 else slice(dy, None)
    w = morph[ys, xs]
    morph[ys, xs] = np.minimum(w, w[::-1, ::-1])
    return morph


def init_extended_morphs(detects, cutoffs, centers, symmetric=True, monotonic=True):
    """Shared tail of init_extended_source (source.py:160-179) for a stack of coadds: sdss symmetry, device
    monotonicity (thresh=.1, float64 like the reference's coadd), noise cut, peak normalisation."""
    detects = np.ascontiguousarray(detects, dtype=np.float64)
    n = len(detects)
    if symmetric:
        for i in range(n):
            _sdss_symmetry(detects[i], centers[i])
    if monotonic and n:
        _native.prox_weighted_monotonic(detects, centers, 0.1)
    ok = np.ones(n, dtype=bool)
    for i in range(n):
        mask = detects[i] > cutoffs[i]
        if mask.sum() == 0:
            ok[i] = False
            continue
        detects[i][~mask] = 0
        detects[i] /= detects[i][int(centers[i][0]), int(centers[i][1])]
    return detects, ok


def init_extended_source(sky_coord, frame, observation, bg_rms, thresh=1., symmetric=True, monotonic=True):
    """Symmetric, monotonic initial morphology + PSF-corrected SED (source.py:139-180)."""
    sed = get_psf_sed(sky_coord, observation, frame)
    if np.any(sed <= 0):
        msg = "Zero or negative SED {} at y={}, x={}".format(sed, *sky_coord)
        if np.all(sed <= 0):
            logger.warning(msg)
        else:
            logger.info(msg)
    morph, bg_cutoff = build_detection_coadd(sed, bg_rms, observation, thresh)
    center = frame.get_pixel(sky_coord)
    morphs, ok = init_extended_morphs(morph[None], [bg_cutoff], [center], symmetric, monotonic)
    if not ok[0]:
        msg = "No flux above threshold={2} for source at y={0} x={1}"
        raise SourceInitError(msg.format(*sky_coord, bg_cutoff))
    return sed, morphs[0]


def init_multicomponent_source(sky_coord, frame, observation, bg_rms, flux_percentiles=None,
                               thresh=1., symmetric=True, monotonic=True):
    """Flux-percentile layered components (source.py:242-291)."""
    if flux_percentiles is None:
        flux_percentiles = [25]
    sed, morph = init_extended_source(sky_coord, frame, observation, bg_rms, thresh, symmetric, monotonic)
    K = len(flux_percentiles) + 1
    Ny, Nx = morph.shape
    morphs = np.zeros((K, Ny, Nx), dtype=morph.dtype)
    morphs[0, :, :] = morph[:, :]
    max_flux = morph.max()
    percentiles_ = np.sort(flux_percentiles)
    last_thresh = 0
    for k in range(1, K):
        flux_thresh = percentiles_[k - 1] * max_flux / 100
        mask_ = morph > flux_thresh
        morphs[k - 1][mask_] = flux_thresh - last_thresh
        morphs[k][mask_] = morph[mask_] - flux_thresh
        last_thresh = flux_thresh
    for k in range(K):
        if np.all(morphs[k] <= 0):
            logger.warning("Zero or negative morphology for component {} at y={}, x={}".format(k, *sky_coord))
        morphs[k] /= morphs[k].max()
    seds = get_best_fit_seds(morphs, frame, observation)
    return seds, morphs


def _centroid_weight(frame):
    """Weight image of psf_weighted_centroid (source.py:389-397)."""
    if frame.psfs is None:
        psf = generate_psf_image(gaussian, (41, 41), amplitude=1, sigma=.9, normalize=False).image
        return psf / psf.max()
    return frame.psfs[0].image


def component_flags(source, component):
    """SCL_* flag word of one component, from its owning source (see include/scarlet_b200.h)."""
    f = 0
    if isinstance(source, PointSource):
        f |= _native.F_CONSTRAINED | _native.F_KSPACE
        if source.symmetric:
            f |= _native.F_SYMMETRIC
        if source.monotonic:
            f |= _native.F_MONOTONIC
    elif isinstance(source, MultiComponentSource):
        # source.py:605-641: one centre per source; its plain Components carry no `.shift` (soft symmetry)
        f |= _native.F_CONSTRAINED | _native.F_SHARED_CENTER
        if source.symmetric:
            f |= _native.F_SYMMETRIC
        if source.monotonic:
            f |= _native.F_MONOTONIC
    if component.fix_sed:
        f |= _native.F_FIX_SED
    if component.fix_morph:
        f |= _native.F_FIX_MORPH
    return f


def constrain_sources(sources, it=0):
    """Run `update()` (source.py:402-440,605-641) once on the device for a list of PointSource-like or
    multi-component sources that share a frame shape, and write the results back (used at construction,
    where it = 0).  Every source is its own one-blend batch entry."""
    if not sources:
        return
    if it != 0:
        raise NotImplementedError("stand-alone update() runs the construction-time pass (it = 0) only; "
                                  "iterations run inside Blend.fit")
    frame = sources[0].frame
    C, Ny, Nx = frame.shape
    n = len(sources)
    comps = [(s, c) for s in sources for c in (s.components if isinstance(s, ComponentTree) else [s])]
    counts = [len(s.components) if isinstance(s, ComponentTree) else 1 for s in sources]
    off = np.concatenate([[0], np.cumsum(counts)])
    if frame.dtype == np.float64:
        # float64 frames (the reference's own loop tests): the same pass by the float64 kernels (csrc/k_f64.cu)
        out = _native.constrain_f64(
            C, Ny, Nx, np.array([c._sed for _, c in comps]), np.array([c._morph for _, c in comps]),
            [component_flags(s, c) for s, c in comps], [s.pixel_center for s, _ in comps],
            [getattr(s, "shift", (0., 0.)) for s, _ in comps], np.asarray(_centroid_weight(frame), dtype=np.float64),
            np.repeat(off[:-1], counts))
    else:
        out = _constrain_f32(frame, sources, comps, counts, off)
    for i, s in enumerate(sources):
        for j, c in enumerate(s.components if isinstance(s, ComponentTree) else [s]):
            c._sed = out["sed"][off[i] + j].astype(frame.dtype)
            c._morph = out["morph"][off[i] + j].astype(frame.dtype)
        k = off[i]
        s.pixel_center = (int(out["center"][k, 0]), int(out["center"][k, 1]))
        if s.symmetric:
            s.shift = (float(out["shift"][k, 0]), float(out["shift"][k, 1]))


def _constrain_f32(frame, sources, comps, counts, off):
    C, Ny, Nx = frame.shape
    n = len(sources)
    cw = np.asarray(_centroid_weight(frame), dtype=np.float32)
    bt = _native.Batch(
        C, Ny, Nx, comp_off=off, sed=np.array([c._sed for _, c in comps]),
        morph=np.array([c._morph for _, c in comps]), cflags=[component_flags(s, c) for s, c in comps],
        center=[s.pixel_center for s, _ in comps], shift=[getattr(s, "shift", (0., 0.)) for s, _ in comps],
        conv=[c.flags.value for _, c in comps], images=np.zeros((n, C, Ny, Nx), np.float32), weights=None,
        kernel_spec=None, ker_idx=-np.ones(n * C), cw=cw[None], cw_idx=np.zeros(n),
        src_first=np.repeat(off[:-1], counts))
    bt.constrain()
    out = bt.download()
    bt.close()
    return out


class RandomSource(Component):
    """Uniform random morphology (source.py:294-320)."""

    def __init__(self, frame, observation=None, **component_kwargs):
        C, Ny, Nx = frame.shape
        morph = np.random.rand(Ny, Nx)
        if observation is None:
            sed = np.random.rand(C)
        else:
            sed = get_best_fit_seds(morph[None], frame, observation)[0]
        super().__init__(frame, sed, morph, **component_kwargs)


class PointSource(Component):
    """Source initialised with the model PSF at one pixel (source.py:323-440)."""

    def __init__(self, frame, sky_coord, observation, symmetric=True, monotonic=True,
                 center_step=5, delay_thresh=10, _defer_update=False, **component_kwargs):
        C, Ny, Nx = frame.shape
        morph = np.zeros((Ny, Nx), observation.images.dtype)
        pixel = frame.get_pixel(sky_coord)
        if frame.psfs is None:
            morph[pixel] = 1
        else:
            psf = frame.psfs[0].image
            assert psf.ndim == 2
            py, px = pixel
            sy, sx = (np.array(psf.shape) - 1) // 2
            cy, cx = (np.array(morph.shape) - 1) // 2
            bottom = int(py - cy - sy) + (Ny >> 1)
            left = int(px - cx - sx) + (Nx >> 1)
            top, right = bottom + psf.shape[0], left + psf.shape[1]
            ys, iys = slice(max(0, bottom), min(Ny, top)), slice(max(0, -bottom), max(Ny - bottom, -top))
            xs, ixs = slice(max(0, left), min(Nx, right)), slice(max(0, -left), max(Nx - left, -right))
            morph[ys, xs] = psf[iys, ixs]
        self.pixel_center = pixel
        pixel = observation.frame.get_pixel(sky_coord)
        sed = observation.images[:, pixel[0], pixel[1]].copy()
        if observation.frame.psfs is not None:
            sed /= observation.frame.psfs.max(axis=(1, 2))
        super().__init__(frame, sed, morph, **component_kwargs)
        self.symmetric = symmetric
        self.monotonic = monotonic
        self.center_step = center_step
        self.delay_thresh = delay_thresh
        if self.symmetric:
            self._centroid_weight = _centroid_weight(self.frame)
        if not _defer_update:
            self.update()

    def update(self):
        """The constraint schedule (source.py:402-440), one pass on the device."""
        it = 0 if self._parent is None else self._parent.it
        constrain_sources([self], it)
        return self


class ExtendedSource(PointSource):
    """Extended source initialised from the detection coadd (source.py:443-492)."""

    def __init__(self, frame, sky_coord, observation, bg_rms, thresh=1, symmetric=True, monotonic=True,
                 center_step=5, delay_thresh=10, _defer_update=False, _init=None, **component_kwargs):
        self.symmetric = symmetric
        self.monotonic = monotonic
        self.coords = sky_coord
        self.pixel_center = frame.get_pixel(sky_coord)
        self.center_step = center_step
        self.delay_thresh = delay_thresh
        if _init is None:
            # the reference always initialises with symmetric=True (source.py:477-478)
            sed, morph = init_extended_source(sky_coord, frame, observation, bg_rms, thresh, True, monotonic)
        else:
            sed, morph = _init
        Component.__init__(self, frame, sed, morph, **component_kwargs)
        if self.symmetric:
            self._centroid_weight = _centroid_weight(self.frame)
        if not _defer_update:
            self.update()


def extended_sources(frame, sky_coords, observation, bg_rms, thresh=1, symmetric=True, monotonic=True,
                     **component_kwargs):
    """Build many ExtendedSources of one frame with batched device calls (same results as
    constructing them one by one; sources whose coadd has no flux above threshold are skipped)."""
    seds, detects, cutoffs, centers = [], [], [], []
    for coord in sky_coords:
        sed = get_psf_sed(coord, observation, frame)
        d, cut = build_detection_coadd(sed, bg_rms, observation, thresh)
        seds.append(sed)
        detects.append(d)
        cutoffs.append(cut)
        centers.append(frame.get_pixel(coord))
    if not seds:
        return []
    morphs, ok = init_extended_morphs(np.array(detects), cutoffs, centers, True, monotonic)
    out = [ExtendedSource(frame, sky_coords[i], observation, bg_rms, thresh, symmetric, monotonic,
                          _defer_update=True, _init=(seds[i], morphs[i]), **component_kwargs)
           for i in range(len(seds)) if ok[i]]
    constrain_sources(out, 0)
    return out


def init_combined_extended_source(sky_coord, frame, observations, bg_rms, obs_idx=0, thresh=1.,
                                  symmetric=True, monotonic=True):
    """SED from every observation (its own WCS and PSF heights), morphology from the detection coadd of
    observations[obs_idx], which must live on the model's pixel grid (source.py:183-240)."""
    try:
        iter(observations)
    except TypeError:
        observations = [observations]
    seds = [get_psf_sed(sky_coord, obs, frame) for obs in observations]
    sed = np.concatenate(seds).flatten()
    if np.any(sed <= 0):
        msg = "Zero or negative SED {} at y={}, x={}".format(sed, *sky_coord)
        (logger.warning if np.all(sed <= 0) else logger.info)(msg)
    detect, bg_cutoff = build_detection_coadd(seds[obs_idx], bg_rms[obs_idx], observations[obs_idx], thresh)
    center = frame.get_pixel(sky_coord)
    morphs, ok = init_extended_morphs(detect[None], [bg_cutoff], [center], symmetric, monotonic)
    if not ok[0]:
        msg = "No flux above threshold={2} for source at y={0} x={1}"
        raise SourceInitError(msg.format(*center, bg_cutoff))
    return sed, morphs[0]


class CombinedExtendedSource(PointSource):
    """Extended source initialised from several observations, e.g. a high- and a low-resolution one
    (source.py:495-535): symmetric defaults to False, the initialisation always symmetrises, and -- like the
    reference -- the constructor does not run `update()`."""

    def __init__(self, frame, sky_coord, observations, bg_rms, obs_idx=0, thresh=1,
                 symmetric=False, monotonic=True, center_step=5, delay_thresh=0, **component_kwargs):
        self.symmetric = symmetric
        self.monotonic = monotonic
        self.coords = sky_coord
        self.pixel_center = frame.get_pixel(sky_coord)
        self.center_step = center_step
        self.delay_thresh = delay_thresh
        sed, morph = init_combined_extended_source(sky_coord, frame, observations, bg_rms, obs_idx,
                                                   thresh, True, monotonic)
        Component.__init__(self, frame, sed, morph, **component_kwargs)
        if self.symmetric:
            self._centroid_weight = _centroid_weight(self.frame)


class MultiComponentSource(ComponentTree):
    """Extended source with vertically layered components (source.py:551-641).  The joint-centre update
    schedule runs on the device: scl_k_source_center measures the centre on the flux-weighted sum of the
    components, the constraint kernel applies symmetry / monotonicity / positivity / normalisation per
    component with that centre (SCL_SHARED_CENTER)."""

    def __init__(self, frame, sky_coord, observation, bg_rms, thresh=1, flux_percentiles=None,
                 symmetric=True, monotonic=True, center_step=5, delay_thresh=10, **component_kwargs):
        self.symmetric = symmetric
        self.monotonic = monotonic
        self.pixel_center = frame.get_pixel(sky_coord)
        self.center_step = center_step
        self.delay_thresh = delay_thresh
        seds, morphs = init_multicomponent_source(sky_coord, frame, observation, bg_rms, flux_percentiles,
                                                  thresh, symmetric, monotonic)
        super().__init__([Component(frame, seds[k], morphs[k]) for k in range(len(seds))])
        if self.symmetric:
            self._centroid_weight = _centroid_weight(self.frame)
        # ensure adherence to constraints (source.py:600-601)
        self.update()

    def update(self):
        it = 0 if self._parent is None else self._parent.it
        constrain_sources([self], it)
        return self
