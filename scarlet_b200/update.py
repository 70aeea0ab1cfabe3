"""Component-level constraint wrappers (the names of scarlet/update.py), dispatching to
the device operators.  Inside `Blend.fit` these are not called one by one: the fused
prox kernel (csrc/k_prox.cu) runs the whole schedule."""
import numpy as np

from . import operator
from . import _native


def positive_sed(component):
    """update.py:13-17 (prox_plus: X[X<0] = 0)"""
    component.sed[component.sed < 0] = 0
    return component


def positive_morph(component):
    """update.py:20-24"""
    component.morph[component.morph < 0] = 0
    return component


def positive(component):
    """update.py:27-32"""
    positive_sed(component)
    positive_morph(component)
    return component


def normalized(component, type='morph_max'):
    """update.py:35-68"""
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


def monotonic(component, pixel_center, use_nearest=False, thresh=0, exact=False, bbox=None):
    """Radial monotonicity from `pixel_center` (update.py:106-156) on the device."""
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
    if bbox is not None:
        raise NotImplementedError("bounding-box restricted updates are not implemented")
    shift = getattr(component, "shift", None)
    morph = component.morph
    operator.prox_uncentered_symmetry(morph, component.step_morph, pixel_center, algorithm, fill, shift, strength)
    component.morph[:] = morph
    return component
