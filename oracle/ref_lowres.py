"""Oracle: the multi-resolution observation (TEST INFRASTRUCTURE ONLY -- never imported by scarlet_b200/).

CPU restatement of /root/reference/scarlet:
  resampling.match_patches                          resampling.py:3-125
  interpolation.sinc_interp / mk_shifter            interpolation.py:302-404
  LowResObservation.match_psfs / build_diffkernel   observation.py:252-343
  LowResObservation.sinc_shift / match / _render / get_loss   observation.py:345-599
  init_combined_extended_source / CombinedExtendedSource      source.py:183-240,495-535

Pinned against outputs of the reference itself (tests/golden/make_golden_lowres.py ->
tests/golden/ref_golden_lowres.npz; tests/test_oracle_lowres.py), unrotated and rotated.  The reference's rotated
branch runs when every low-resolution pixel lies inside the high-resolution frame (reshape at
observation.py:486-489) and the operator cache holds no real=True shifter of the same FFT shape (mk_shifter's cache
key ignores `real`, interpolation.py:319; tests/golden/repro_rotated_lowres.py shows both outcomes); the oracle
restates the working (cold-cache) behaviour.  Quirks kept on purpose are marked "quirk".

`grad` (what autograd derives for `get_loss`): `_render` is linear in the model, so its adjoint is the transpose
of its matrix; the matrix is built by rendering unit models through the restated `_render`.
"""
import copy
import numpy as np
from . import ref_fft, ref_prox
from . import ref_blend as rb


def match_patches(shape_hr, shape_lr, wcs_hr, wcs_lr, isrot=False):
    """resampling.py:3-125 with perimeter='overlap'.  isrot: every low-resolution pixel is listed (flattened grid,
    resampling.py:53-60); otherwise rows and columns separately."""
    Ny_hr, Nx_hr = shape_hr[-2:]
    Ny_lr, Nx_lr = shape_lr[-2:]
    y_hr, x_hr = np.arange(Ny_hr), np.arange(Nx_hr)
    if isrot:
        Y_lr, X_lr = np.indices((Ny_lr, Nx_lr))
        X_lr, Y_lr = X_lr.flatten(), Y_lr.flatten()
    else:
        Y_lr, X_lr = np.arange(Ny_lr), np.arange(Nx_lr)
    ra_lr, dec_lr = wcs_lr.all_pix2world(X_lr, Y_lr, 0, ra_dec_order=True)
    ra_hr, dec_hr = wcs_hr.all_pix2world(x_hr, y_hr, 0, ra_dec_order=True)
    X_hr, Y_hr = wcs_hr.all_world2pix(ra_lr, dec_lr, 0, ra_dec_order=True)
    x_lr, y_lr = wcs_lr.all_world2pix(ra_hr, dec_hr, 0, ra_dec_order=True)
    # quirk: per-column and per-row conditions multiplied elementwise (frames must be square), bounds N+1
    over_lr = (X_hr >= 0) * (X_hr < Nx_hr + 1) * (Y_hr >= 0) * (Y_hr < Ny_hr + 1)
    over_hr = (x_lr >= 0) * (x_lr < Nx_lr + 1) * (y_lr >= 0) * (y_lr < Ny_lr + 1)
    coordhr_hr = (y_hr[over_hr == 1], x_hr[over_hr == 1])
    if np.sum(over_lr) == 0:
        raise rb.SourceInitError("no overlap")
    coordlr_lr = (Y_lr[over_lr == 1], X_lr[over_lr == 1])
    coordlr_hr = (Y_hr[over_lr == 1], X_hr[over_lr == 1])
    return coordlr_lr, coordlr_hr, coordhr_hr


def sinc_interp(images, coord_hr, coord_lr, angle=None, padding=3):
    """interpolation.py:342-404."""
    y_hr, x_hr = coord_hr
    y_lr, x_lr = coord_lr
    hy = np.abs(y_lr[1] - y_lr[0])
    hx = np.abs(x_lr[1] - x_lr[0])
    assert hy != 0 and hx != 0
    if angle is None:
        # quirk: the image enters transposed (interpolation.py:370)
        return np.array([np.dot(np.dot(np.sinc((y_lr[np.newaxis, :] - y_hr[:, np.newaxis]) / hy), image.T),
                                np.sinc((x_lr[:, np.newaxis] - x_hr[np.newaxis, :]) / hx)) for image in images])
    # rotated grids (interpolation.py:374-404): shift every image to each high-resolution row in Fourier space, then
    # the two 1-D sinc sums
    cos, sin = angle
    shape = ref_fft.fft_shape(images.shape, images.shape, padding=padding, axes=[1, 2])
    X_fft = ref_fft.forward(images, shape, (1, 2))
    shifter_y, shifter_x = mk_shifter(shape)
    shift_y = shifter_y[np.newaxis, :] ** (-y_hr[:, np.newaxis] * cos / hy)
    shift_x = shifter_x[np.newaxis, :] ** (-y_hr[:, np.newaxis] * sin / hx)
    result_fft = X_fft[:, np.newaxis, :, :] * shift_y[np.newaxis, :, :, np.newaxis]
    result_fft = result_fft * shift_x[np.newaxis, :, np.newaxis, :]
    result_shape = (result_fft.shape[0], result_fft.shape[1], images.shape[1], images.shape[2])
    result_shift = ref_fft.inverse(result_fft, shape, result_shape, (2, 3))
    shy = np.sinc((y_lr[np.newaxis, :] + x_hr[:, np.newaxis] * sin) / hy)
    shx = np.sinc((x_lr[np.newaxis, :] - x_hr[:, np.newaxis] * cos) / hx)
    result_y = (result_shift[:, :, np.newaxis, :, :] * shy[np.newaxis, np.newaxis, :, :, np.newaxis]).sum(axis=-2)
    return (result_y * shx[np.newaxis, np.newaxis, :, :]).sum(axis=-1)


def mk_shifter(shape, real=False):
    """interpolation.py:302-340 (without the cache)."""
    freq_x = np.fft.rfftfreq(shape[1])
    freq_y = np.fft.rfftfreq(shape[0]) if real else np.fft.fftfreq(shape[0])
    return np.exp(-1j * 2 * np.pi * freq_y), np.exp(-1j * 2 * np.pi * freq_x)


class LowResObservation(rb.Observation):
    def __init__(self, images, wcs=None, psfs=None, weights=None, channels=None):
        assert wcs is not None and psfs is not None
        super().__init__(images, psfs=psfs, weights=weights, channels=channels)
        self.frame.wcs = wcs
        self._dense = None

    # observation.py:252-317
    def match_psfs(self, psf_hr, wcs_hr, angle=None):
        psf_lr = self.frame.psfs
        ny_hr, nx_hr = psf_hr.shape
        _, ny_lr, nx_lr = psf_lr.shape
        w_hr, w_lr = wcs_hr.deepcopy(), self.frame.wcs.deepcopy()
        w_hr.wcs.crval = 0., 0.
        w_hr.wcs.crpix = ny_hr / 2. + 1, nx_hr / 2. + 1
        w_lr.wcs.crval = 0., 0.
        w_lr.wcs.crpix = ny_lr / 2. + 1, nx_lr / 2. + 1
        pl_lr, pl_hr, over_hr = match_patches(psf_hr.shape, psf_lr.shape[1:], w_hr, w_lr)
        npsf_y = np.max(over_hr[0]) - np.min(over_hr[0]) + 1
        npsf_x = np.max(over_hr[1]) - np.min(over_hr[1]) + 1
        # quirk: the row slice ends at the COLUMN maximum (observation.py:305)
        psf_valid = psf_lr[:, pl_lr[0].min():pl_lr[1].max() + 1, pl_lr[1].min():pl_lr[1].max() + 1]
        psf_match_lr = sinc_interp(psf_valid, over_hr, pl_hr, angle=angle)
        psf_match_hr = psf_hr[over_hr[0].min():over_hr[0].max() + 1,
                              over_hr[1].min():over_hr[1].max() + 1].reshape(npsf_y, npsf_x)
        assert psf_match_lr[0].shape == psf_match_hr.shape
        psf_match_hr = psf_match_hr / np.sum(psf_match_hr)
        psf_match_lr = psf_match_lr / np.sum(psf_match_lr)     # quirk: one normalisation over all bands
        return psf_match_hr[np.newaxis, :], psf_match_lr

    # observation.py:319-343
    def build_diffkernel(self, model_frame, angle=None):
        target = model_frame.psfs[0, :, :]
        new_target, observed = self.match_psfs(target, model_frame.wcs, angle)
        return ref_fft.match_psfs(observed, new_target)

    # observation.py:345-403: one transformed axis (unrotated frames) or both (rotated frames)
    def sinc_shift(self, imgs, shifts, axes, sum_axis):
        axes = tuple(np.array(axes) - 1)
        sum_axis -= 1
        if len(axes) == 1:
            ax = axes[0]
            n_ax = self._fft_shape[ax]
            spec = ref_fft.forward(imgs, [n_ax], (ax + 1,))
            shifter = mk_shifter(self._fft_shape, real=True)
            sh = shifter[ax][np.newaxis, :] ** shifts[ax][:, np.newaxis]
            # "convolution by sinc" (observation.py:381-385): the slice bounds are n//4 and (-n)//4 of the HALF spectrum
            sh[:, sh.shape[1] // 4:-sh.shape[1] // 4] = 0
            if ax == 0:
                spec = spec[:, np.newaxis, :, :] * sh[np.newaxis, :, :, np.newaxis]
            else:
                spec = spec[:, np.newaxis, :, :] * sh[np.newaxis, :, np.newaxis, :]
            out_shape = spec.shape[:2] + tuple(imgs.shape[1:])
            return ref_fft.inverse(spec, [n_ax], out_shape, (ax + 2,))
        # both axes (observation.py:371-403 with axes = (1, 2)): a full transform along y, a real one along x
        fshape = np.array(self._fft_shape)
        spec = ref_fft.forward(imgs, fshape, (1, 2))
        shifter = mk_shifter(self._fft_shape)
        shishift = [shifter[ax][np.newaxis, :] ** np.asarray(shifts[ax])[:, np.newaxis] for ax in axes]
        n = shishift[sum_axis].shape[1]
        shishift[sum_axis][:, n // 4:-n // 4] = 0
        spec = spec[:, np.newaxis, :, :] * shishift[0][np.newaxis, :, :, np.newaxis]
        spec = spec * shishift[1][np.newaxis, :, np.newaxis, :]
        out_shape = tuple(spec.shape[:2]) + tuple(fshape)
        return ref_fft.inverse(spec, fshape, out_shape, (2, 3))

    # observation.py:405-521
    def match(self, model_frame):
        if self.frame.dtype != model_frame.dtype:
            self.frame.dtype = model_frame.dtype
            self.images = self.images.astype(model_frame.dtype)
            if isinstance(self.weights, np.ndarray):
                self.weights = self.weights.astype(model_frame.dtype)
            self.frame.psfs = self.frame.psfs.astype(model_frame.dtype)
        self.band_slice = slice(None)
        if self.frame.channels is not model_frame.channels:
            bmin = list(model_frame.channels).index(self.frame.channels[0])
            bmax = list(model_frame.channels).index(self.frame.channels[-1])
            self.band_slice = slice(bmin, bmax + 1)
        ma, sa = model_frame.wcs.wcs.pc, self.frame.wcs.wcs.pc
        model_pix = np.sqrt(np.abs(ma[0, 0]) * np.abs(ma[1, 1] - ma[0, 1] * ma[1, 0]))
        self_pix = np.sqrt(np.abs(sa[0, 0]) * np.abs(sa[1, 1] - sa[0, 1] * sa[1, 0]))
        sv = np.sum(sa, axis=0)[:2] / self_pix
        mv = np.sum(ma, axis=0)[:2] / model_pix
        sv = sv / np.sum(sv ** 2) ** .5
        mv = mv / np.sum(mv ** 2) ** .5
        self.sin_rot = sv[0] * mv[1] - sv[1] * mv[0]          # np.cross of two 2-vectors (observation.py:444)
        self.cos_rot = np.dot(sv, mv)
        self.isrot = np.abs(self.sin_rot) ** 2 > np.finfo(float).eps
        if not self.isrot:
            self.sin_rot, self.cos_rot, angle = 0, 1, None
        else:
            angle = (self.cos_rot, self.sin_rot)
        coord_lr, coord_hr, _ = match_patches(model_frame.shape, self.frame.shape, model_frame.wcs, self.frame.wcs,
                                              isrot=self.isrot)
        self.lr_shape = (np.max(coord_lr[0]) - np.min(coord_lr[0]) + 1, np.max(coord_lr[1]) - np.min(coord_lr[1]) + 1)
        self._coord_lr, self._coord_hr = coord_lr, coord_hr
        center_y = (np.max(coord_hr[0]) - np.min(coord_hr[0]) + 1) / 2
        center_x = (np.max(coord_hr[1]) - np.min(coord_hr[1]) + 1) / 2
        diff_psf = self.build_diffkernel(model_frame, angle)
        self.small_axis = self.frame.shape[2] <= self.frame.shape[1]
        self._fft_shape = ref_fft.fft_shape(model_frame.psfs.shape, model_frame.shape, padding=3, axes=[-2, -1],
                                            use_max=True)
        diff_psf = ref_fft.pad(diff_psf, self._fft_shape, axes=(1, 2))
        if self.isrot:
            # observation.py:483-503: the rotated grid described by its un-rotated rows and columns (every low-resolution
            # pixel must lie inside the model frame for the reshape)
            Y_unrot = ((coord_hr[0] - center_y) * self.cos_rot + (coord_hr[1] - center_x) * self.sin_rot).reshape(self.lr_shape)
            X_unrot = ((coord_hr[1] - center_x) * self.cos_rot - (coord_hr[0] - center_y) * self.sin_rot).reshape(self.lr_shape)
            self.Y_unrot, self.X_unrot = Y_unrot[:, 0], X_unrot[0, :]
            if self.small_axis:
                self.shifts = [self.Y_unrot * self.cos_rot, self.Y_unrot * self.sin_rot]
                self.other_shifts = [-self.sin_rot * self.X_unrot, self.cos_rot * self.X_unrot]
            else:
                self.shifts = [-self.sin_rot * self.X_unrot, self.cos_rot * self.X_unrot]
                self.other_shifts = [self.Y_unrot * self.cos_rot, self.Y_unrot * self.sin_rot]
            axes = (1, 2)
        else:
            axes = [int(not self.small_axis) + 1]
            self.shifts = np.array(coord_hr, dtype=np.float64)
            self.shifts[0] -= center_y
            self.shifts[1] -= center_x
            self.other_shifts = np.copy(self.shifts)
        op = self.sinc_shift(diff_psf, self.shifts, axes, int(not self.small_axis) + 1)
        self._resconv_op = np.array(op, dtype=self.frame.dtype) * (self_pix / model_pix) ** 2
        self._resconv_op = self._resconv_op.reshape(*self._resconv_op.shape[:2], -1)
        self._dense = None
        return self

    # observation.py:524-558
    def _render(self, model):
        model_ = ref_fft.pad(model[self.band_slice, :, :], self._fft_shape, axes=(-2, -1))
        axes = (1, 2) if self.isrot else [int(self.small_axis) + 1]
        conv = self.sinc_shift(model_, self.other_shifts, axes, int(self.small_axis) + 1)
        conv = conv.reshape(*conv.shape[:2], -1)
        C = self.frame.shape[0]
        if self.small_axis:
            return np.array([(conv[c] @ self._resconv_op[c].T).T for c in range(C)], dtype=self.frame.dtype)[:, :, ::-1]
        return np.array([conv[c] @ self._resconv_op[c].T for c in range(C)], dtype=self.frame.dtype)[:, ::-1, :]

    render_lr = _render

    def _window(self):
        y0, x0 = int(np.min(self._coord_lr[0])), int(np.min(self._coord_lr[1]))
        y1, x1 = int(np.max(self._coord_lr[0])), int(np.max(self._coord_lr[1]))
        return slice(y0, y1 + 1), slice(x0, x1 + 1)

    # observation.py:560-578 (display)
    def render(self, model):
        img = np.zeros(self.frame.shape)
        ys, xs = self._window()
        img[:, ys, xs] = self._render(model)
        return img

    # observation.py:580-599
    def get_loss(self, model):
        ys, xs = self._window()
        return 0.5 * np.sum((self.weights * (self._render(model) - self.images[:, ys, xs])) ** 2)

    def dense_operator(self, model_shape):
        """[C_lr, n_lr, Ny*Nx] matrix of `_render` (linear in the model)."""
        if self._dense is None:
            _, Ny, Nx = model_shape
            Cl = self.frame.shape[0]
            n = int(np.prod(self.lr_shape))
            L = np.zeros((Cl, n, Ny * Nx), np.float64)
            e = np.zeros(model_shape, np.float64)
            band = e[self.band_slice].reshape(Cl, -1)
            keep = self.frame.dtype
            self.frame.dtype = np.float64
            for p in range(Ny * Nx):
                band[:, p] = 1
                L[:, :, p] = self._render(e).reshape(Cl, n)
                band[:, p] = 0
            self.frame.dtype = keep
            self._dense = L
        return self._dense

    def grad(self, model):
        """d get_loss / d model: transpose of `_render` applied to W^2 (render - images)."""
        ys, xs = self._window()
        r = self.weights * self.weights * (self._render(model) - self.images[:, ys, xs])
        L = self.dense_operator(model.shape)
        Cl = self.frame.shape[0]
        g = np.zeros(model.shape, dtype=model.dtype)
        g[self.band_slice] = np.einsum("cq,cqp->cp", r.reshape(Cl, -1).astype(np.float64), L).reshape(
            (Cl,) + model.shape[1:]).astype(model.dtype)
        return g


# ---------------------------------------------------------------- source.py:183-240,495-535
def get_pixel(frame, sky_coord):
    """Frame.get_pixel (observation.py:85-99)."""
    wcs = getattr(frame, "wcs", None)
    if wcs is not None:
        coord = wcs.wcs_world2pix(sky_coord[0], sky_coord[1], 0)
        return (int(coord[0].item()), int(coord[1].item()))
    return tuple(int(c) for c in sky_coord)


def get_psf_sed(sky_coord, obs, frame):
    """source.py:21-72 through the observation's own WCS."""
    pixel = get_pixel(obs.frame, sky_coord)
    sed = obs.images[:, pixel[0], pixel[1]].copy()
    if obs.frame.psfs is not None:
        sed /= obs.frame.psfs.max(axis=(1, 2))
    if frame.psfs is not None:
        sed = sed * frame.psfs[0].max()
    return sed


def init_combined_extended_source(sky_coord, frame, observations, bg_rms, obs_idx=0, thresh=1.,
                                  symmetric=True, monotonic=True):
    """source.py:183-240"""
    seds = [get_psf_sed(sky_coord, obs, frame) for obs in observations]
    sed = np.concatenate(seds).flatten()
    morph, bg_cutoff = rb.build_detection_coadd(seds[obs_idx], bg_rms[obs_idx], observations[obs_idx], thresh)
    center = get_pixel(frame, sky_coord)
    if symmetric:
        morph = ref_prox.prox_uncentered_symmetry(morph, 0, center=center, algorithm="sdss")
    if monotonic:
        morph = rb._monotonic_cached(morph.shape, center, .1)(morph).reshape(morph.shape)
    mask = morph > bg_cutoff
    if mask.sum() == 0:
        raise rb.SourceInitError("no flux above threshold")
    morph[~mask] = 0
    morph /= morph[int(center[0]), int(center[1])]
    return sed, morph


class CombinedExtendedSource(rb.PointSource):
    """source.py:495-535 (symmetric defaults to False, the initialisation always symmetrises)."""

    def __init__(self, frame, sky_coord, observations, bg_rms, obs_idx=0, thresh=1, symmetric=False, monotonic=True,
                 **kw):
        self.symmetric, self.monotonic = symmetric, monotonic
        self.pixel_center = get_pixel(frame, sky_coord)
        sed, morph = init_combined_extended_source(sky_coord, frame, observations, bg_rms, obs_idx, thresh, True,
                                                   monotonic)
        rb.Component.__init__(self, frame, sed, morph, **kw)
