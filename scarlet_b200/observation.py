"""Frame / Observation: the data side of the model (mirrors scarlet/observation.py:13-239).

`Observation.match` builds the PSF difference kernels once (host, fft.match_psfs);
`render` and `get_loss` run the device convolution.  Inside `Blend.fit` none of these
Python methods is called: the fused kernel consumes `images`, `weights` and the kernel
spectrum directly (blend.py of this package).
"""
import logging

import numpy as np

from . import fft
from . import _native

logger = logging.getLogger("scarlet.observation")


class Frame():
    """Spatial and spectral characteristics of the data (observation.py:13-99)."""

    def __init__(self, shape, wcs=None, psfs=None, channels=None, dtype=np.float32):
        assert len(shape) == 3
        self._shape = tuple(shape)
        self.wcs = wcs
        if psfs is None:
            logger.warning('No PSFs specified. Possible, but dangerous!')
        else:
            msg = 'PSFs need to have shape (1,Ny,Nx) for Blend and (B,Ny,Nx) for Observation'
            assert len(psfs) == 1 or len(psfs) == shape[0], msg
            if not isinstance(psfs, fft.Fourier):
                psfs = fft.Fourier(np.array(psfs))
            if not np.allclose(psfs.sum(axis=(1, 2)), 1):
                logger.warning('PSFs not normalized. Normalizing now..')
                psfs.normalize(axes=(1, 2))
            if dtype != psfs.image.dtype:
                logger.warning("Dtypes of PSFs and Frame different. Casting PSFs to {}".format(dtype))
                psfs.update_dtype(dtype)
        self._psfs = psfs
        assert channels is None or len(channels) == shape[0]
        self.channels = channels
        self.dtype = dtype

    @property
    def C(self):
        return self._shape[0]

    @property
    def Ny(self):
        return self._shape[1]

    @property
    def Nx(self):
        return self._shape[2]

    @property
    def shape(self):
        return self._shape

    @property
    def psfs(self):
        return self._psfs

    def get_pixel(self, sky_coord):
        if self.wcs is not None:
            if self.wcs.naxis == 3:
                coord = self.wcs.wcs_world2pix(sky_coord[0], sky_coord[1], 0, 0)
            elif self.wcs.naxis == 2:
                coord = self.wcs.wcs_world2pix(sky_coord[0], sky_coord[1], 0)
            else:
                raise ValueError("Invalid number of wcs dimensions: {0}".format(self.wcs.naxis))
            return (int(coord[0].item()), int(coord[1].item()))
        return tuple(int(coord) for coord in sky_coord)


class Observation():
    """Data and metadata for a single set of observations (observation.py:101-239)."""

    def __init__(self, images, psfs=None, weights=None, wcs=None, channels=None, padding=10):
        self.frame = Frame(images.shape, wcs=wcs, psfs=psfs, channels=channels, dtype=images.dtype)
        self.images = np.array(images)
        self.weights = np.array(weights) if weights is not None else 1
        self._padding = padding
        self._band_slice = slice(None)
        self._diff_kernels = None
        self._spec = None

    def match(self, model_frame):
        if self.frame.dtype != model_frame.dtype:
            msg = "Dtypes of model and observation different. Casting observation to {}"
            logger.warning(msg.format(model_frame.dtype))
            self.frame.dtype = model_frame.dtype
            self.images = self.images.astype(model_frame.dtype)
            if type(self.weights) is np.ndarray:
                self.weights = self.weights.astype(model_frame.dtype)
            if self.frame._psfs is not None:
                self.frame.psfs.update_dtype(model_frame.dtype)
        self._band_slice = slice(None)
        if self.frame.channels is not model_frame.channels:
            assert self.frame.channels is not None and model_frame.channels is not None
            bmin = list(model_frame.channels).index(self.frame.channels[0])
            bmax = list(model_frame.channels).index(self.frame.channels[-1])
            self._band_slice = slice(bmin, bmax + 1)
        self._diff_kernels = None
        self._spec = None
        if self.frame.psfs is not model_frame.psfs:
            assert self.frame.psfs is not None and model_frame.psfs is not None
            p_obs, p_mod = self.frame.psfs.image, model_frame.psfs.image
            if model_frame.dtype == np.float64:
                # float64 frames (the reference's own loop tests): the k-space division in float64 on the host,
                # exactly fft.match_psfs (fft.py:282-301); the float64 loop convolves with the real-space kernels
                self._diff_kernels = fft.match_psfs(self.frame.psfs, model_frame.psfs)
            elif p_obs.ndim == 3 and p_mod.ndim == 3 and p_mod.shape[0] == 1 and len(self.frame.shape) == 3:
                # the usual case (one model PSF, one observed PSF per band): kernels and their spectra in one
                # device pass (csrc/k_psf.cu)
                diff, spec = _native.match_psfs(p_obs[None], p_mod, frame_shape=self.frame.shape[1:])
                self._diff_kernels = fft.Fourier(diff[0].astype(p_obs.dtype, copy=False))
                self._spec = spec[0]
            else:
                self._diff_kernels = fft.match_psfs(self.frame.psfs, model_frame.psfs)
        return self

    def kernel_spectrum(self):
        """complex64 [C,128,65] spectra of the truncated difference kernels for the device, or None."""
        if self._diff_kernels is None:
            return None
        if self._spec is None:
            _, Ny, Nx = self.frame.shape
            self._spec = _native.kernel_spectrum(self._diff_kernels.image, Ny, Nx)
        return self._spec

    def _convolve(self, model):
        C = model.shape[0]
        _, ky, kx = self._diff_kernels.image.shape
        if model.dtype == np.float64:
            return _native.convolve_f64(model[None], np.broadcast_to(self._diff_kernels.image, (C, ky, kx)))[0]
        out = _native.convolve(np.asarray(model, dtype=np.float32)[None], self.kernel_spectrum(), np.arange(C),
                               ker_half=_native.kernel_half(ky, kx, model.shape[1], model.shape[2]))[0]
        return out.astype(model.dtype, copy=False)

    def render(self, model):
        """Convolve a model to the observation frame (observation.py:203-220)."""
        model_ = model[self._band_slice, :, :]
        if self._diff_kernels is not None:
            model_ = self._convolve(model_)
        return model_

    def get_loss(self, model):
        """0.5 sum (W (render(model) - images))^2 (observation.py:222-239)."""
        model = self.render(model)
        return 0.5 * np.sum((self.weights * (model - self.images)) ** 2)


class LowResObservation(Observation):
    """Observation at a coarser pixel scale than the model (observation.py:242-599).

    `match` does the geometry on the host, once: pixel correspondence through the two WCS (resampling.match_patches),
    the difference kernel between the sinc-resampled observation PSFs and the model PSF, the difference kernel
    sinc-shifted to every low-resolution row (`_resconv_op`, as the reference stores it) -- and then folds the
    reference's whole `_render` (pad, FFT along x, per-column phase shift + its "sinc" mask, inverse FFT, contraction
    with `_resconv_op`, final flip) into one matrix per band, `_lr_op[c]`: [n_lr, Ny*Nx].  `_render`, `get_loss` and
    the loop apply that matrix on the device (scl_lowres_render_f32 / scl_k_lowres).  Unrotated frames must be square
    (the reference multiplies per-row and per-column masks, resampling.py:90-92); rotated frames (`_match_rotated`) must
    lie inside the model frame, as in the reference (observation.py:486-489).
    """

    def __init__(self, images, wcs=None, psfs=None, weights=None, channels=None, padding=3, operator='exact'):
        assert wcs is not None, "WCS is necessary for LowResObservation"
        assert psfs is not None, "PSFs are necessary for LowResObservation"
        assert operator in ['exact', 'bilinear', 'SVD']
        super().__init__(images, wcs=wcs, psfs=psfs, weights=weights, channels=channels, padding=padding)
        self._lr_op = None

    @staticmethod
    def _centred_wcs(wcs, ny, nx):
        w = wcs.deepcopy()
        if w.naxis == 2:
            w.wcs.crval = 0., 0.
            w.wcs.crpix = ny / 2. + 1, nx / 2. + 1
        elif w.naxis == 3:
            w.wcs.crval = 0., 0., 0.
            w.wcs.crpix = ny / 2. + 1, nx / 2. + 1, 0.
        return w

    def match_psfs(self, psf_hr, wcs_hr, angle):
        """Model PSF and observation PSFs on one grid at the model's resolution (observation.py:252-317)."""
        from . import resampling, interpolation
        psf_lr = self.frame.psfs.image
        ny_hr, nx_hr = psf_hr.shape
        _, ny_lr, nx_lr = psf_lr.shape
        w_hr = self._centred_wcs(wcs_hr, ny_hr, nx_hr)
        w_lr = self._centred_wcs(self.frame.wcs, ny_lr, nx_lr)
        at_lr, at_hr, over_hr = resampling.match_patches(psf_hr.shape, psf_lr.shape[1:], w_hr, w_lr, isrot=False, psf=True)
        # the reference ends the row slice at the column maximum (observation.py:305)
        lo_y, lo_x, hi = at_lr[0].min(), at_lr[1].min(), at_lr[1].max() + 1
        resampled = interpolation.sinc_interp(psf_lr[:, lo_y:hi, lo_x:hi], over_hr, at_hr, angle=angle)
        target = psf_hr[over_hr[0].min():over_hr[0].max() + 1, over_hr[1].min():over_hr[1].max() + 1]
        assert resampled[0].shape == target.shape
        # one normalisation over all bands together, like the reference (observation.py:315-316)
        return (target / target.sum())[None], resampled / resampled.sum()

    def build_diffkernel(self, model_frame, angle):
        """observation.py:319-343"""
        target, observed = self.match_psfs(model_frame.psfs.image[0], model_frame.wcs, angle)
        return fft.match_psfs(fft.Fourier(observed), fft.Fourier(target))

    def _shift_kernels(self, shifts, n):
        """Real circular kernels k_s[n] of the reference's 1-D "sinc shift" (observation.py:345-403): phase ramp
        to the power s on the rfft frequencies, the band [m//4, -m//4) of that HALF spectrum (m = n//2+1) zeroed
        (observation.py:381-385 as written), inverse real transform."""
        from . import interpolation
        ramp = interpolation.mk_shifter((n, n), real=True)[0]
        spec = ramp[None, :] ** np.asarray(shifts, dtype=np.float64)[:, None]
        m = spec.shape[1]
        spec[:, m // 4:-m // 4] = 0
        return np.fft.irfft(spec, n, axis=1)

    def match(self, model_frame):
        from . import resampling
        if self.frame.dtype != model_frame.dtype:
            msg = "Dtypes of model and observation different. Casting observation to {}"
            logger.warning(msg.format(model_frame.dtype))
            self.frame.dtype = model_frame.dtype
            self.images = self.images.astype(model_frame.dtype)
            if type(self.weights) is np.ndarray:
                self.weights = self.weights.astype(model_frame.dtype)
            self.frame.psfs.update_dtype(model_frame.dtype)
        self._band_slice = slice(None)
        if self.frame.channels is not model_frame.channels:
            bmin = list(model_frame.channels).index(self.frame.channels[0])
            bmax = list(model_frame.channels).index(self.frame.channels[-1])
            self._band_slice = slice(bmin, bmax + 1)

        def affine(w):
            try:
                return np.asarray(w.wcs.pc)
            except AttributeError:
                return np.asarray(w.cd)
        ma, sa = affine(model_frame.wcs), affine(self.frame.wcs)
        model_pix = np.sqrt(np.abs(ma[0, 0]) * np.abs(ma[1, 1] - ma[0, 1] * ma[1, 0]))
        self_pix = np.sqrt(np.abs(sa[0, 0]) * np.abs(sa[1, 1] - sa[0, 1] * sa[1, 0]))
        sv, mv = np.sum(sa, axis=0)[:2] / self_pix, np.sum(ma, axis=0)[:2] / model_pix
        sv, mv = sv / np.sum(sv ** 2) ** .5, mv / np.sum(mv ** 2) ** .5
        self.sin_rot, self.cos_rot = sv[0] * mv[1] - sv[1] * mv[0], np.dot(sv, mv)
        self.isrot = bool(np.abs(self.sin_rot) ** 2 > np.finfo(float).eps)
        if not self.isrot:
            self.sin_rot, self.cos_rot = 0, 1
        angle = (self.cos_rot, self.sin_rot) if self.isrot else None
        coord_lr, coord_hr, _ = resampling.match_patches(model_frame.shape, self.frame.shape, model_frame.wcs,
                                                        self.frame.wcs, isrot=self.isrot)
        self.lr_shape = (int(coord_lr[0].max() - coord_lr[0].min() + 1), int(coord_lr[1].max() - coord_lr[1].min() + 1))
        self._coord_lr, self._coord_hr = coord_lr, coord_hr
        self.frame_coord = (np.arange(model_frame.Ny), np.arange(model_frame.Nx))
        center_y = (coord_hr[0].max() - coord_hr[0].min() + 1) / 2
        center_x = (coord_hr[1].max() - coord_hr[1].min() + 1) / 2
        diff_psf = self.build_diffkernel(model_frame, angle)
        self.small_axis = self.frame.Nx <= self.frame.Ny          # always True on square frames
        if self.isrot:
            return self._match_rotated(model_frame, diff_psf, coord_hr, center_y, center_x, (self_pix / model_pix) ** 2)
        self._fft_shape = fft._get_fft_shape(model_frame.psfs, np.zeros(model_frame.shape), padding=3,
                                             axes=[-2, -1], max=True)
        Fy, Fx = self._fft_shape
        self.shifts = np.array(coord_hr, dtype=np.float64)
        self.shifts[0] -= center_y
        self.shifts[1] -= center_x
        self.other_shifts = np.copy(self.shifts)
        # _resconv_op[c, j] = difference kernel shifted along y to low-resolution row j  (observation.py:516-519)
        dpad = fft._pad(np.asarray(diff_psf.image, dtype=np.float64), self._fft_shape, axes=(1, 2))
        ky = self._shift_kernels(self.shifts[0], Fy)                             # [n_y, Fy]
        rows = (np.arange(Fy)[:, None] - np.arange(Fy)[None, :]) % Fy
        resconv = np.einsum("juw,cwv->cjuv", ky[:, rows], dpad) * (self_pix / model_pix) ** 2
        self._resconv_op = resconv.reshape(resconv.shape[0], resconv.shape[1], -1).astype(self.frame.dtype)
        # fold `_render` into one matrix per band:
        #   out[c, j, i'] = sum_{u,v} (model shifted along x to column i)[u, v] * resconv[c, j, u, v],  i = n_x-1-i'
        Ny, Nx = model_frame.Ny, model_frame.Nx
        py, px = (Fy - Ny + 1) // 2, (Fx - Nx + 1) // 2                          # fft._pad offsets
        kx = self._shift_kernels(self.other_shifts[1], Fx)                       # [n_x, Fx]
        cols = (np.arange(Fx)[:, None] - (np.arange(Nx)[None, :] + px)) % Fx
        KX = kx[::-1][:, cols]                                                   # [n_x (flipped), Fx, Nx]
        # [c, j, y, v] . [i, v, x] over v -> [c, j, y, i, x] (one BLAS call) -> [c, j, i, y, x]
        op = np.tensordot(resconv[:, :, py:py + Ny, :], KX, axes=([3], [1])).transpose(0, 1, 3, 2, 4)
        C_lr, n_y, n_x = op.shape[:3]
        assert (n_y, n_x) == self.lr_shape
        self._lr_op = np.ascontiguousarray(op.reshape(C_lr, n_y * n_x, Ny * Nx), dtype=np.float32)
        return self

    def _shift2(self, imgs, shifts, masked_axis):
        """The reference's two-axis "sinc shift" (observation.py:345-403 with axes = (1, 2)) of a stack imgs [n, Fy, Fx]
        already on the FFT box: for every entry k of the shift lists the stack moved by (shifts[0][k], shifts[1][k])
        pixels with Fourier phase ramps (full transform along y, real transform along x), the ramp of `masked_axis`
        low-passed by zeroing its entries [m//4, -m//4).  Returns [n, K, Fy, Fx]."""
        from . import interpolation
        Fy, Fx = self._fft_shape
        spec = np.fft.rfftn(np.fft.ifftshift(imgs, axes=(1, 2)), axes=(1, 2))
        ramps = interpolation.mk_shifter((Fy, Fx))
        phase = [ramps[ax][None, :] ** np.asarray(shifts[ax], dtype=np.float64)[:, None] for ax in (0, 1)]
        m = phase[masked_axis].shape[1]
        phase[masked_axis][:, m // 4:-m // 4] = 0
        spec = spec[:, None] * phase[0][None, :, :, None] * phase[1][None, :, None, :]
        return np.fft.fftshift(np.fft.irfftn(spec, (Fy, Fx), axes=(2, 3)), axes=(2, 3))

    def _match_rotated(self, model_frame, diff_psf, coord_hr, center_y, center_x, flux_scale):
        """The rotated branch of `match` (observation.py:483-503): the low-resolution grid is described by its
        un-rotated rows Y and columns X; row j sees the difference kernel moved by (Y_j cos, Y_j sin), column i the
        model moved by (-X_i sin, X_i cos), and pixel (j, i) is the product of the two summed over the FFT box
        (`_render`, observation.py:524-558).  Both factors are circular convolutions on the box, so the whole map is one
        matrix per band,  op[c, (j, i'), (y, x)] = (R_cj correlated with h_i)[y, x],  h_i = the moved unit model pixel:
        built with one FFT product per (band, row, column) and handed to the device like the unrotated operator."""
        if not self.small_axis:
            raise NotImplementedError("rotated low-resolution frames wider than tall are not supported")
        n_y, n_x = self.lr_shape
        if len(coord_hr[0]) != n_y * n_x:
            raise NotImplementedError("a rotated low-resolution frame must lie entirely inside the model frame "
                                      "(the reference reshapes its pixel list to the frame, observation.py:486-489)")
        Yu = ((coord_hr[0] - center_y) * self.cos_rot + (coord_hr[1] - center_x) * self.sin_rot).reshape(n_y, n_x)[:, 0]
        Xu = ((coord_hr[1] - center_x) * self.cos_rot - (coord_hr[0] - center_y) * self.sin_rot).reshape(n_y, n_x)[0, :]
        self.Y_unrot, self.X_unrot = Yu, Xu
        self.shifts = [Yu * self.cos_rot, Yu * self.sin_rot]
        self.other_shifts = [-self.sin_rot * Xu, self.cos_rot * Xu]
        self._fft_shape = fft._get_fft_shape(model_frame.psfs, np.zeros(model_frame.shape), padding=3,
                                             axes=[-2, -1], max=True)
        Fy, Fx = self._fft_shape
        Ny, Nx = model_frame.Ny, model_frame.Nx
        dpad = fft._pad(np.asarray(diff_psf.image, dtype=np.float64), self._fft_shape, axes=(1, 2))
        R = self._shift2(dpad, self.shifts, 0) * flux_scale                         # [C, n_y, Fy, Fx]
        self._resconv_op = R.reshape(R.shape[0], n_y, -1).astype(self.frame.dtype)
        unit = np.zeros((1, Ny, Nx))
        unit[0, 0, 0] = 1
        h = self._shift2(fft._pad(unit, self._fft_shape, axes=(1, 2)), self.other_shifts, 1)[0]     # [n_x, Fy, Fx]
        hf = np.conj(np.fft.rfft2(h[::-1]))                                         # column order flipped (i' = n_x-1-i)
        C_lr = R.shape[0]
        op = np.empty((C_lr, n_y, n_x, Ny, Nx), np.float32)
        for c in range(C_lr):
            rf = np.fft.rfft2(R[c])                                                 # [n_y, Fy, Fx/2+1]
            op[c] = np.fft.irfft2(rf[:, None] * hf[None], (Fy, Fx))[:, :, :Ny, :Nx]
        self._lr_op = np.ascontiguousarray(op.reshape(C_lr, n_y * n_x, Ny * Nx))
        return self

    def _window(self):
        y0, x0 = int(self._coord_lr[0].min()), int(self._coord_lr[1].min())
        return slice(y0, y0 + self.lr_shape[0]), slice(x0, x0 + self.lr_shape[1])

    def _render(self, model):
        """The model resampled and convolved into the low-resolution window (observation.py:524-558), on the device."""
        m = np.asarray(model[self._band_slice], dtype=np.float32)
        out = _native.lowres_render(m.reshape(1, m.shape[0], -1), self._lr_op)[0]
        return out.reshape((m.shape[0],) + self.lr_shape).astype(self.frame.dtype, copy=False)

    def render(self, model):
        """For display: the window inside an otherwise empty observation frame (observation.py:560-578)."""
        img = np.zeros(self.frame.shape)
        ys, xs = self._window()
        img[:, ys, xs] = self._render(model)
        return img

    def get_loss(self, model):
        """observation.py:580-599"""
        ys, xs = self._window()
        return 0.5 * np.sum((self.weights * (self._render(model) - self.images[:, ys, xs])) ** 2)
