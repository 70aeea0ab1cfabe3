"""Component-level constraint wrappers (the names of scarlet/update.py), dispatching to
the device operators.  Inside `Blend.fit` these are not called one by one: the fused
prox kernel (csrc/k_prox.cu) runs the whole schedule."""
import numpy as np

from . import operator
from . import _native
from . import schedule


def positive_sed(component):
    """update.py:13-17 (prox_plus: X[X<0] = 0)"""
    if schedule.recording() is not None:
        schedule.record("positive_sed")
        return component
    component.sed[component.sed < 0] = 0
    return component


def positive_morph(component):
    """update.py:20-24"""
    if schedule.recording() is not None:
        schedule.record("positive_morph")
        return component
    component.morph[component.morph < 0] = 0
    return component


def positive(component):
    """update.py:27-32"""
    positive_sed(component)
    positive_morph(component)
    return component


def normalized(component, type='morph_max'):
    """update.py:35-68"""
    if schedule.recording() is not None:
        kinds = {'sed': 'norm_sed', 'morph': 'norm_morph', 'morph_max': 'norm_morph_max'}
        if type.lower() not in kinds:
            raise ValueError("Unrecognized normalization '{0}'".format(type))
        schedule.record(kinds[type.lower()])
        return component
    t = type.lower()
    if t == 'sed':
        norm = component.sed.sum()
        component.sed[:] = component.sed / norm
        component.morph[:] = component.morph * norm
    elif t == 'morph':
        norm = component.morph.sum()
        component.sed[:] = component.sed * norm
        component.morph[:] = component.morph / norm
    elif t == 'morph_max':
        norm = component.morph.max()
        component.sed[:] = component.sed * norm
        component.morph[:] = component.morph / norm
    else:
        raise ValueError("Unrecognized normalization '{0}'".format(type))
    return component


def _on_device(morph, fn):
    """Run `fn` on a float32 C-contiguous copy of `morph` and write the result back in morph's dtype."""
    x = np.ascontiguousarray(morph, dtype=np.float32)
    out = fn(x)
    if x is not morph:
        morph[...] = x
    return out


def sparse_l0(component, thresh):
    """L0 sparsity of the morphology: pixels with |morph| < thresh * step_morph are zeroed (update.py:71-75,
    proxmin prox_hard)."""
    if schedule.recording() is not None:
        schedule.record("sparse_l0", thresh)
        return component
    _on_device(component.morph, lambda x: _native.prox_sparse(x, thresh * component.step_morph, 0))
    return component


def sparse_l1(component, thresh):
    """L1 sparsity: soft threshold of the morphology by thresh * step_morph (update.py:78-82, proxmin prox_soft)."""
    if schedule.recording() is not None:
        schedule.record("sparse_l1", thresh)
        return component
    _on_device(component.morph, lambda x: _native.prox_sparse(x, thresh * component.step_morph, 1))
    return component


def threshold(component):
    """Zero the pixels below the noise cut of measurement.threshold and keep the box of the rest in
    `component.bboxes["thresh"]` (update.py:85-103)."""
    if schedule.recording() is not None:
        schedule.record("threshold")        # the box of the rest is a post-fit product (scl_batch_products)
        return component
    from .bbox import Box
    _, _, bounds = _on_device(component.morph, lambda x: _native.threshold(x[None], apply=True))
    b = bounds[0]
    if b[0] < 0:
        raise ValueError("zero-size array to reduction operation minimum which has no identity")
    if not hasattr(component, "bboxes"):
        component.bboxes = {}
    component.bboxes["thresh"] = Box.from_bounds(int(b[0]), int(b[1]), int(b[2]), int(b[3]))
    return component


def translation(component, direction=1, kernel=None, padding=3):
    """Shift the morphology by `direction` times the component's sub-pixel `shift` (update.py:159-167; like the
    reference, the resampling always uses the lanczos kernel whatever `kernel` is passed)."""
    if schedule.recording() is not None:
        raise NotImplementedError("update.translation cannot run inside the CUDA loop")
    from . import interpolation
    dy, dx = component.shift
    component.morph[:] = interpolation.fft_resample(component.morph, dy * direction, dx * direction)
    return component


def monotonic(component, pixel_center, use_nearest=False, thresh=0, exact=False, bbox=None):
    """Radial monotonicity from `pixel_center` (update.py:106-156) on the device."""
    if schedule.recording() is not None:
        if use_nearest or exact or bbox is not None:
            raise NotImplementedError("inside the CUDA loop update.monotonic runs the weighted operator on the whole frame")
        schedule.live_center(component, pixel_center, "update.monotonic")
        schedule.record("monotonic", thresh)
        return component
    if exact:
        raise NotImplementedError("Exact monotonicity is not currently supported")
    if bbox is not None:
        raise NotImplementedError("bounding-box restricted updates are not implemented")
    prox = operator.prox_strict_monotonic(component.shape[-2:], use_nearest=use_nearest, thresh=thresh,
                                          center=pixel_center)
    morph = component.morph.copy()
    prox(morph, component.step_morph)
    component.morph[:] = morph
    return component


def symmetric(component, pixel_center, algorithm="kspace", bbox=None, fill=None, strength=.5):
    """Symmetry about `pixel_center` (update.py:170-199) on the device."""
    if schedule.recording() is not None:
        if bbox is not None or fill is not None:
            raise NotImplementedError("inside the CUDA loop update.symmetric runs on the whole frame without fill")
        schedule.live_center(component, pixel_center, "update.symmetric")
        if algorithm == "kspace":
            # operator.py:337-339: without a `.shift` the k-space request falls back to soft symmetry, strength 1
            if hasattr(component, "shift"):
                schedule.record("sym_kspace")
            else:
                schedule.record("sym_soft", 1.0)
        elif algorithm == "soft":
            schedule.record("sym_soft", strength)
        elif algorithm == "sdss":
            schedule.record("sym_sdss")
        else:
            raise ValueError("algorithm must be one of 'soft', 'sdss', 'kspace', recieved '{0}''".format(algorithm))
        return component
    if bbox is not None:
        raise NotImplementedError("bounding-box restricted updates are not implemented")
    shift = getattr(component, "shift", None)
    morph = component.morph
    operator.prox_uncentered_symmetry(morph, component.step_morph, pixel_center, algorithm, fill, shift, strength)
    component.morph[:] = morph
    return component
