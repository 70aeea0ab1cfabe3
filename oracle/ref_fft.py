"""Oracle: padding / cropping / FFT-shape conventions and the k-space
convolution + PSF matching of the reference (TEST INFRASTRUCTURE ONLY).

Follows /root/reference/scarlet/fft.py:
  centered()        fft.py:7-35
  pad()             fft.py:38-65
  next_fast_len()   scipy.fftpack.helper.next_fast_len as called at fft.py:99,104
                    (5-smooth sizes; third-party, restated from its published
                    definition: smallest 2^a 3^b 5^c >= target)
  fft_shape()       fft.py:68-106
  forward()/inverse()  Fourier.fft fft.py:193-211 / Fourier.from_fft fft.py:138-181
  kspace_op()       fft.py:264-279
  match_psfs()      fft.py:282-301
  convolve()        fft.py:304-317
"""
import numpy as np


def centered(arr, newshape):
    newshape = np.asarray(newshape)
    cur = np.array(arr.shape)
    if not np.all(newshape <= cur):
        raise ValueError("arr must be larger than newshape in both dimensions, "
                         "received {0}, and {1}".format(arr.shape, newshape))
    start = (cur - newshape + 1) // 2
    sl = tuple(slice(int(s), int(s + n)) for s, n in zip(start, newshape))
    return arr[sl]


def pad(arr, newshape, axes=None):
    if axes is None:
        axes = tuple(range(arr.ndim))
    widths = [(0, 0)] * arr.ndim
    for a, ax in enumerate(axes):
        dS = int(newshape[a]) - arr.shape[ax]
        left = (dS + 1) // 2
        widths[ax] = (left, dS - left)
    return np.pad(arr, widths, mode="constant")


def next_fast_len(target):
    target = int(target)
    if target <= 6:
        return target
    best = None
    p5 = 1
    while p5 < 2 * target:
        p35 = p5
        while p35 < 2 * target:
            n = p35
            while n < target:
                n *= 2
            if best is None or n < best:
                best = n
            p35 *= 3
        p5 *= 5
    return best


def fft_shape(shape1, shape2, padding=3, axes=None, use_max=False):
    shape1 = np.asarray(shape1)
    shape2 = np.asarray(shape2)
    if len(shape1) != len(shape2):
        raise ValueError("img1 and img2 must have the same number of dimensions")
    if axes is None:
        axes = range(len(shape1))
    out = []
    for ax in axes:
        s = max(shape1[ax], shape2[ax]) if use_max else shape1[ax] + shape2[ax]
        out.append(next_fast_len(s + padding))
    while out[-1] % 2 != 0:
        out[-1] = next_fast_len(out[-1] + 1)
    return out


def forward(image, shape, axes):
    """rfftn(ifftshift(pad(image)))  (fft.py:209-210)"""
    padded = pad(image, shape, axes)
    return np.fft.rfftn(np.fft.ifftshift(padded, axes), axes=axes)


def inverse(spec, shape, image_shape, axes):
    """centered(fftshift(irfftn(spec)))  (fft.py:173-178)"""
    img = np.fft.irfftn(spec, shape, axes=axes)
    img = np.fft.fftshift(img, axes=axes)
    return centered(img, image_shape)


def kspace_op(img1, img2, padding, op, out_shape, axes):
    shape = fft_shape(img1.shape, img2.shape, padding, axes)
    spec = op(forward(img1, shape, axes), forward(img2, shape, axes))
    return inverse(spec, shape, out_shape, axes)


def match_psfs(psf1, psf2, padding=3, axes=(-2, -1)):
    """Difference kernel psf1 / psf2 in k-space (fft.py:282-301).

    The output is cropped to the shape of whichever stack has the larger
    FIRST dimension (band count), exactly as fft.py:297-300 compares shape[0].
    """
    shape = psf2.shape if psf1.shape[0] < psf2.shape[0] else psf1.shape
    axes = tuple(a % psf1.ndim for a in axes)
    return kspace_op(psf1, psf2, padding, np.divide, shape, axes)


def convolve(img, kernel, padding=3, axes=(-2, -1)):
    axes = tuple(a % img.ndim for a in axes)
    return kspace_op(img, kernel, padding, np.multiply, img.shape, axes)


def correlate(img, kernel, padding=3, axes=(-2, -1)):
    """Exact transpose of `convolve` w.r.t. img: same pad/crop, conj(kernel
    spectrum).  This is what autograd derives for fft.py:304-317
    (SURVEY Appendix A.3)."""
    axes = tuple(a % img.ndim for a in axes)
    shape = fft_shape(img.shape, kernel.shape, padding, axes)
    spec = forward(img, shape, axes) * np.conj(forward(kernel, shape, axes))
    return inverse(spec, shape, img.shape, axes)
