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
        out = _native.convolve(np.asarray(model, dtype=np.float32)[None], self.kernel_spectrum(), np.arange(C))[0]
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
    """Multi-resolution observation (observation.py:242-599): SURVEY 8(a) row A14, not built yet."""

    def __init__(self, *args, **kwargs):
        raise NotImplementedError("LowResObservation (multi-resolution rendering) is not implemented in scarlet_b200 yet")
