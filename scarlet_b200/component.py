"""Component / ComponentTree: the parameter containers of a blend (the names and behaviour of scarlet/component.py).

They hold host numpy copies of SED and morphology; `Blend.fit` packs them into device buffers, runs the CUDA loop
and writes the results back (`_sed` / `_morph` are rebound, as in the reference where every step rebinds them,
blend.py:92-95).  Nothing here computes on pixels inside the loop.
"""
import enum
import logging

import numpy as np

logger = logging.getLogger("scarlet.component")


class BlendFlag(enum.Flag):
    """Status bits of a component (component.py:13-36); the device keeps the same bit values
    (`scl_batch_products`, include/scarlet_b200.h)."""
    NONE = 0
    SED_NOT_CONVERGED = 1
    MORPH_NOT_CONVERGED = 2
    EDGE_PIXELS = 4
    NO_VALID_PIXELS = 8


class Prior():
    """Differentiable prior hook (component.py:39-67).  A Python callback cannot run inside the CUDA loop: `Blend.fit`
    accepts the priors whose gradient the device can evaluate itself (`AffinePrior`) and rejects any other."""

    def __init__(self, grad_func, L_func):
        self._grad_func, self._L_func = grad_func, L_func
        self.sed_grad = self.morph_grad = 0

    def compute_grad(self, component):
        state = (component._sed, component._morph)
        self.sed_grad, self.morph_grad = self._grad_func(*state)
        self.L_sed, self.L_morph = self._L_func(*state)


class AffinePrior(Prior):
    """A prior whose gradient is affine in the parameter it acts on,

        grad_sed = a_sed * sed + b_sed            grad_morph = a_morph * morph + b_morph,

    which covers the quadratic penalties lambda/2 |x - x0|^2 (a = lambda, b = -lambda x0: Gaussian priors around a
    template SED or morphology, Tikhonov damping with x0 = 0) and linear ones (a = 0: a constant pull, e.g. an L1
    flux penalty on non-negative parameters).  `b_sed` [C] and `b_morph` [Ny, Nx] may be scalars.  The Lipschitz
    constants of the two gradients default to |a|.  Additive to the reference (which only has the callback form,
    component.py:39-67); the same object is a valid reference-style Prior -- `compute_grad` evaluates the same
    expressions on the host -- and `Blend.fit` ships (a, b, L) to the device, where Component.backward_prior
    (component.py:177-187) runs inside the gradient step of the constraint kernel."""

    def __init__(self, a_sed=0., b_sed=0., a_morph=0., b_morph=0., L_sed=None, L_morph=None):
        self.a_sed, self.a_morph = float(a_sed), float(a_morph)
        self.b_sed, self.b_morph = b_sed, b_morph
        self._L = (abs(self.a_sed) if L_sed is None else float(L_sed),
                   abs(self.a_morph) if L_morph is None else float(L_morph))
        Prior.__init__(self, lambda sed, morph: (self.a_sed * sed + self.b_sed, self.a_morph * morph + self.b_morph),
                       lambda sed, morph: self._L)


class _Node:
    """Position of a component or subtree inside its parent tree."""
    _index = None
    _parent = None

    @property
    def coord(self):
        """Indices leading from the root tree to this node (None for a node outside any tree)."""
        if self._index is None:
            return None
        above = self._parent.coord if self._parent._index is not None else ()
        return tuple(above) + (self._index,)

    def _attach(self, parent, index):
        self._parent, self._index = parent, index


def _inverse_or_none(obj, name):
    try:
        return 1 / getattr(obj, name)
    except AttributeError:
        return None


class Component(_Node):
    """One SED x morphology term of the scene model (component.py:70-210)."""

    def __init__(self, frame, sed, morph, prior=None, fix_sed=False, fix_morph=False):
        self._frame = frame
        dtype = frame.dtype
        self._sed = np.array(sed, dtype=dtype)
        self._morph = np.array(morph, dtype=dtype)
        self._last_sed = np.zeros_like(self._sed)
        self._last_morph = np.zeros_like(self._morph)
        self.prior = prior
        self.fix_sed, self.fix_morph = fix_sed, fix_morph
        self.sed_grad = self.morph_grad = 0
        self.L_sed = self.L_morph = 1
        self.flags = BlendFlag.SED_NOT_CONVERGED | BlendFlag.MORPH_NOT_CONVERGED

    frame = property(lambda self: self._frame)
    shape = property(lambda self: self._frame.shape)
    sed = property(lambda self: self._sed)
    morph = property(lambda self: self._morph)
    step_sed = property(lambda self: _inverse_or_none(self, "L_sed"))
    step_morph = property(lambda self: _inverse_or_none(self, "L_morph"))

    def get_model(self, sed=None, morph=None):
        """sed[c] * morph[y, x] as a (C, Ny, Nx) cube; both arguments or neither (component.py:148-170)."""
        if (sed is None) != (morph is None):
            raise ValueError("You need to supply `sed` and `morph` or neither")
        if sed is None:
            sed, morph = self._sed, self._morph
        return sed[:, None, None] * morph[None, :, :]

    def get_flux(self):
        """Total flux per band (component.py:172-175)."""
        return self.morph.sum() * self.sed

    def backward_prior(self):
        """Add the prior's gradients and Lipschitz terms to the free parameters (component.py:177-190)."""
        if self.prior is None:
            return
        self.prior.compute_grad(self)
        for fixed, what in ((self.fix_morph, "morph"), (self.fix_sed, "sed")):
            if not fixed:
                setattr(self, what + "_grad", getattr(self, what + "_grad") + getattr(self.prior, what + "_grad"))
                setattr(self, "L_" + what, getattr(self, "L_" + what) + getattr(self.prior, "L_" + what))

    def update(self):
        """Constraint hook; sources override it (component.py:192-196)."""
        return self


class ComponentTree(_Node):
    """Sources and sub-trees of a scene, addressable by position (component.py:213-408)."""

    def __init__(self, components):
        nodes = tuple(components) if hasattr(components, "__iter__") else (components,)
        for node in nodes:
            if not isinstance(node, (Component, ComponentTree)):
                raise NotImplementedError("argument needs to be list of Components or ComponentTrees")
        self._tree = nodes
        for i, node in enumerate(nodes):
            assert node.frame is self.frame, "All components need to share the same Frame"
            node._attach(self, i)
        self._components = None

    # ---- structure
    @property
    def components(self):
        """The leaf components in tree order, each object once."""
        if self._components is None:
            seen, flat = set(), []
            for node in self._tree:
                for leaf in (node.components if isinstance(node, ComponentTree) else (node,)):
                    if id(leaf) not in seen:
                        seen.add(id(leaf))
                        flat.append(leaf)
            self._components = tuple(flat)
        return self._components

    n_components = property(lambda self: len(self.components))
    K = property(lambda self: len(self.components))
    sources = property(lambda self: self._tree)
    n_sources = property(lambda self: len(self._tree))
    frame = property(lambda self: self._tree[0].frame)

    def __iter__(self):
        return iter(self._tree)

    def __len__(self):
        return len(self._tree)

    def __getitem__(self, coord):
        if isinstance(coord, int):
            return self._tree[coord]
        if isinstance(coord, (tuple, list)):
            node = self._tree[coord[0]]
            return node[coord[1:]] if len(coord) > 1 else node
        raise NotImplementedError("coord needs to be index or list of indices")

    def __iadd__(self, other):
        if isinstance(other, ComponentTree):
            grown = self._tree + other._tree
        elif isinstance(other, Component):
            grown = self._tree + (other,)
        else:
            raise NotImplementedError("argument needs to be Component or ComponentTree")
        other._attach(self, self.n_sources)
        self._tree = grown
        self._components = None
        return self

    # ---- model
    def get_model(self, seds=None, morphs=None):
        """Sum of the component models, optionally for given parameters (component.py:321-347)."""
        model = np.zeros(self.frame.shape, dtype=self.frame.dtype)
        given = seds is not None and morphs is not None
        for k, comp in enumerate(self.components):
            model = model + (comp.get_model(seds[k], morphs[k]) if given else comp.get_model())
        return model

    def get_flux(self):
        """Flux per band summed over the components (component.py:349-357)."""
        total = None
        for comp in self.components:
            total = comp.get_flux() if total is None else total + comp.get_flux()
        return total

    def update(self):
        for node in self._tree:
            node.update()
