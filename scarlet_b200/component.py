"""Component / ComponentTree: the parameter containers (mirrors scarlet/component.py).

They hold host numpy copies of SED and morphology; `Blend.fit` packs them into device
buffers, runs the CUDA loop and writes the results back (`_sed`/`_morph` are replaced,
as in the reference where every step rebinds them, blend.py:92-95).
"""
from enum import Flag, auto
import logging

import numpy as np

logger = logging.getLogger("scarlet.component")


class BlendFlag(Flag):
    """Flags that can be set by scarlet (component.py:13-36)."""
    NONE = 0
    SED_NOT_CONVERGED = auto()
    MORPH_NOT_CONVERGED = auto()
    EDGE_PIXELS = auto()
    NO_VALID_PIXELS = auto()


class Prior():
    """Differentiable prior hook (component.py:39-67).  Kept for API parity; `Blend.fit`
    rejects components with a prior (a Python callback cannot run inside the CUDA loop)."""

    def __init__(self, grad_func, L_func):
        self._grad_func = grad_func
        self._L_func = L_func
        self.sed_grad = 0
        self.morph_grad = 0

    def compute_grad(self, component):
        self.sed_grad, self.morph_grad = self._grad_func(component._sed, component._morph)
        self.L_sed, self.L_morph = self._L_func(component._sed, component._morph)


class Component():
    """A single component in a blend (component.py:70-210)."""

    def __init__(self, frame, sed, morph, prior=None, fix_sed=False, fix_morph=False):
        self._frame = frame
        self._sed = np.array(sed, dtype=self._frame.dtype)
        self._morph = np.array(morph, dtype=self._frame.dtype)
        self.sed_grad = 0
        self.morph_grad = 0
        self.prior = prior
        self.L_sed = 1
        self.L_morph = 1
        self.flags = BlendFlag.SED_NOT_CONVERGED | BlendFlag.MORPH_NOT_CONVERGED
        self._last_sed = np.zeros(self._sed.shape, dtype=self._sed.dtype)
        self._last_morph = np.zeros(self._morph.shape, dtype=self._morph.dtype)
        self._index = None
        self._parent = None
        self.fix_sed = fix_sed
        self.fix_morph = fix_morph

    @property
    def shape(self):
        return self._frame.shape

    @property
    def coord(self):
        if self._index is not None:
            if self._parent._index is not None:
                return tuple(self._parent.coord) + (self._index,)
            return (self._index,)

    @property
    def frame(self):
        return self._frame

    @property
    def sed(self):
        return self._sed

    @property
    def morph(self):
        return self._morph

    def get_model(self, sed=None, morph=None):
        if sed is not None and morph is not None:
            return sed[:, None, None] * morph[None, :, :]
        elif sed is None and morph is None:
            return self._sed[:, None, None] * self._morph[None, :, :]
        raise ValueError("You need to supply `sed` and `morph` or neither")

    def get_flux(self):
        return self.morph.sum() * self.sed

    def backward_prior(self):
        if self.prior is not None:
            self.prior.compute_grad(self)
            if not self.fix_morph:
                self.morph_grad += self.prior.morph_grad
                self.L_morph += self.prior.L_morph
            if not self.fix_sed:
                self.sed_grad += self.prior.sed_grad
                self.L_sed += self.prior.L_sed

    def update(self):
        return self

    @property
    def step_morph(self):
        try:
            return 1 / self.L_morph
        except AttributeError:
            return None

    @property
    def step_sed(self):
        try:
            return 1 / self.L_sed
        except AttributeError:
            return None


class ComponentTree():
    """Hierarchical collection of Components (component.py:213-408)."""

    def __init__(self, components):
        if not hasattr(components, "__iter__"):
            components = (components,)
        self._tree = tuple(components)
        self._index = None
        self._parent = None
        for i, c in enumerate(self._tree):
            if not isinstance(c, ComponentTree) and not isinstance(c, Component):
                raise NotImplementedError("argument needs to be list of Components or ComponentTrees")
            assert c.frame is self.frame, "All components need to share the same Frame"
            c._index = i
            c._parent = self
        self._components = None

    @property
    def components(self):
        if self._components is None:
            components = []
            for c in self._tree:
                _c = c.components if isinstance(c, ComponentTree) else [c]
                for __c in _c:
                    if not any(__c is o for o in components):
                        components.append(__c)
            self._components = tuple(components)
        return self._components

    @property
    def n_components(self):
        return len(self.components)

    @property
    def K(self):
        return self.n_components

    @property
    def frame(self):
        return self._tree[0].frame

    @property
    def sources(self):
        return self._tree

    @property
    def n_sources(self):
        return len(self._tree)

    @property
    def coord(self):
        if self._index is not None:
            if self._parent._index is not None:
                return tuple(self._parent.coord) + (self._index,)
            return (self._index,)

    def get_model(self, seds=None, morphs=None):
        model = np.zeros(self.frame.shape, dtype=self.frame.dtype)
        for k in range(self.K):
            if seds is not None and morphs is not None:
                model = model + self.components[k].get_model(seds[k], morphs[k])
            else:
                model = model + self.components[k].get_model()
        return model

    def get_flux(self):
        for k, component in enumerate(self.components):
            if k == 0:
                model = component.get_flux()
            else:
                model += component.get_flux()
        return model

    def update(self):
        for node in self._tree:
            node.update()

    def __iadd__(self, c):
        c_index = self.n_sources
        if isinstance(c, ComponentTree):
            self._tree = self._tree + c._tree
        elif isinstance(c, Component):
            self._tree = self._tree + (c,)
        else:
            raise NotImplementedError("argument needs to be Component or ComponentTree")
        c._index = c_index
        c._parent = self
        self._components = None
        return self

    def __getitem__(self, coord):
        if isinstance(coord, (tuple, list)):
            if len(coord) > 1:
                return self._tree[coord[0]].__getitem__(coord[1:])
            return self._tree[coord[0]]
        elif isinstance(coord, int):
            return self._tree[coord]
        raise NotImplementedError("coord needs to be index or list of indices")

    def __iter__(self):
        return iter(self._tree)

    def __len__(self):
        return len(self._tree)
