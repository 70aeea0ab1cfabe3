"""Padding/cropping conventions, the `Fourier` image container and PSF matching
(host-side SETUP code mirroring scarlet/fft.py), plus `convolve`, which runs on the
device.

On the hot path the per-iteration FFT convolution lives in the fused CUDA kernel
(csrc/k_forward.cu); what remains here is what the reference runs once per blend:
`match_psfs` (fft.py:282-301) and the container semantics user code relies on
(`.image`, `.shape`, indexing, normalisation).
"""
import numpy as np

from . import _native


def _centered(arr, newshape):
    """Centre `newshape` portion of arr, odd-in-even goes centre-right (fft.py:7-35)."""
    newshape = np.asarray(newshape)
    currshape = np.array(arr.shape)
    if not np.all(newshape <= currshape):
        msg = "arr must be larger than newshape in both dimensions, received {0}, and {1}"
        raise ValueError(msg.format(arr.shape, newshape))
    start = (currshape - newshape + 1) // 2
    return arr[tuple(slice(int(s), int(s + n)) for s, n in zip(start, newshape))]


def _pad(arr, newshape, axes=None):
    """Zero-pad to newshape with the fftshift convention (fft.py:38-65)."""
    if axes is None:
        axes = range(arr.ndim)
    try:
        len(axes)
    except TypeError:
        axes = [axes]
    widths = [(0, 0)] * arr.ndim
    for a, ax in enumerate(axes):
        dS = int(newshape[a]) - arr.shape[ax]
        left = (dS + 1) // 2
        widths[ax] = (left, dS - left)
    return np.pad(arr, widths, mode="constant")


def _get_fft_shape(img1, img2, padding=3, axes=None, max=False):
    """Fast (5-smooth) FFT lengths per axis, last one even (fft.py:68-106)."""
    shape1, shape2 = np.asarray(img1.shape), np.asarray(img2.shape)
    if len(shape1) != len(shape2):
        msg = "img1 and img2 must have the same number of dimensions, but got {0} and {1}"
        raise ValueError(msg.format(len(shape1), len(shape2)))
    if axes is None:
        axes = range(len(shape1))
    try:
        len(axes)
    except TypeError:
        axes = [axes]
    shape = []
    for ax in axes:
        s = np.max([shape1[ax], shape2[ax]]) if max else shape1[ax] + shape2[ax]
        shape.append(_native.next_fast_len(s + padding))
    while shape[-1] % 2 != 0:
        shape[-1] = _native.next_fast_len(shape[-1] + 1)
    return shape


class Fourier(object):
    """An image that remembers its Fourier transforms (fft.py:109-261)."""

    def __init__(self, image, image_fft=None):
        self._fft = {} if image_fft is None else image_fft
        self._image = image

    @staticmethod
    def from_fft(image_fft, fft_shape, image_shape, axes=None):
        if axes is None:
            axes = range(len(image_fft))
        image = np.fft.irfftn(image_fft, fft_shape, axes=axes)
        image = np.fft.fftshift(image, axes=axes)
        image = _centered(image, image_shape)
        key = (tuple(fft_shape), tuple(axes), tuple(range(len(image_shape))))
        return Fourier(image, {key: image_fft})

    @property
    def image(self):
        return self._image

    @property
    def shape(self):
        return self._image.shape

    def fft(self, fft_shape, axes):
        try:
            iter(axes)
        except TypeError:
            axes = (axes,)
        key = (tuple(fft_shape), tuple(axes), tuple(range(self._image.ndim)))
        if key not in self._fft:
            if len(fft_shape) != len(axes):
                msg = "fft_shape self.axes must have the same number of dimensions, got {0}, {1}"
                raise ValueError(msg.format(fft_shape, axes))
            image = _pad(self._image, fft_shape, axes)
            self._fft[key] = np.fft.rfftn(np.fft.ifftshift(image, axes), axes=axes)
        return self._fft[key]

    def __len__(self):
        return len(self._image)

    def normalize(self, axes=None):
        if axes is not None:
            idx = [slice(None)] * self._image.ndim
            for ax in axes:
                idx[ax] = None
        else:
            idx = [None] * self._image.ndim
        norm = 1 / self._image.sum(axis=axes)
        self._image *= norm[tuple(idx)]
        self._fft = {}

    def update_dtype(self, dtype):
        if self._image.dtype != dtype:
            self._image = self._image.astype(dtype)
            self._fft = {}

    def sum(self, axis=None):
        return self._image.sum(axis)

    def max(self, axis=None):
        return self._image.max(axis=axis)

    def __getitem__(self, index):
        return Fourier(self._image[index])


def _kspace_operation(image1, image2, padding, op, shape, axes):
    if len(image1.shape) != len(image2.shape):
        msg = "Both images must have the same number of axes, got {0} and {1}"
        raise Exception(msg.format(len(image1.shape), len(image2.shape)))
    axes = tuple(a % len(image1.shape) for a in axes)
    fft_shape = _get_fft_shape(image1.image, image2.image, padding, axes)
    spec = op(image1.fft(fft_shape, axes), image2.fft(fft_shape, axes))
    return Fourier.from_fft(spec, fft_shape, shape, axes)


def match_psfs(psf1, psf2, padding=3, axes=(-2, -1)):
    """Difference kernel psf1 / psf2 in k-space (fft.py:282-301); once per blend, on the host."""
    shape = psf2.shape if psf1.shape[0] < psf2.shape[0] else psf1.shape
    return _kspace_operation(psf1, psf2, padding, np.divide, shape, axes)


def convolve(image1, image2, padding=3, axes=(-2, -1)):
    """image1 (*) image2 over the last two axes, on the device (fft.py:304-317).

    image1: `Fourier` of a 2-D image or a (C, Ny, Nx) cube; image2: `Fourier` of a kernel
    with the same number of axes (one kernel per band, or one broadcast to all bands).
    """
    img, ker = np.asarray(image1.image), np.asarray(image2.image)
    if img.ndim != ker.ndim:
        msg = "Both images must have the same number of axes, got {0} and {1}"
        raise Exception(msg.format(img.ndim, ker.ndim))
    nd = img.ndim
    if tuple(a % nd for a in axes) != (nd - 2, nd - 1) or nd not in (2, 3):
        raise NotImplementedError("device convolution runs over the last two axes of 2-D or 3-D arrays")
    cube = img.reshape((-1,) + img.shape[-2:])
    kers = ker.reshape((-1,) + ker.shape[-2:])
    C, Ny, Nx = cube.shape
    spec = _native.kernel_spectrum(kers, Ny, Nx)
    idx = np.arange(C) if len(kers) == C else np.zeros(C, dtype=np.int32)
    out = _native.convolve(cube[None].astype(np.float32), spec, idx,
                           ker_half=_native.kernel_half(kers.shape[1], kers.shape[2], Ny, Nx))[0]
    return Fourier(out.reshape(img.shape).astype(img.dtype, copy=False))
