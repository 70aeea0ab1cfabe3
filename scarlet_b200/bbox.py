"""Bounding boxes (the names of scarlet/bbox.py).  `Box` is a host container; `trim` and `flux_at_edge` scan the
image on the device (csrc/k_products.cu, `scl_trim_f32`), any number of images per call through `trim_many`."""
import numpy as np

from . import _native


class Box(object):
    """(y, x) corner + height + width of a rectangle of pixels (bbox.py:4-167).

    A box with a non-positive edge is empty: `yx0` is None, its bounds are None and its slices select nothing.
    """

    def __init__(self, yx0, height, width):
        empty = width <= 0 or height <= 0
        self.yx0 = None if empty else yx0
        self.height = 0 if empty else height
        self.width = 0 if empty else width

    @staticmethod
    def from_bounds(bottom, top, left, right):
        """Inclusive bounds -> Box (bbox.py:29-49)."""
        return Box((bottom, left), top + 1 - bottom, right + 1 - left)

    @property
    def is_empty(self):
        return self.width == 0 or self.height == 0

    def _bound(self, axis, far):
        if self.is_empty:
            return None
        return self.yx0[axis] + ((self.height, self.width)[axis] - 1 if far else 0)

    bottom = property(lambda self: self._bound(0, False))
    top = property(lambda self: self._bound(0, True))
    left = property(lambda self: self._bound(1, False))
    right = property(lambda self: self._bound(1, True))

    @property
    def slices(self):
        if self.is_empty:
            return slice(0, 0), slice(0, 0)
        return slice(self.bottom, self.top + 1), slice(self.left, self.right + 1)

    @property
    def shape(self):
        return (self.height, self.width)

    def __or__(self, other):
        """Smallest box holding both (bbox.py:107-124)."""
        return Box.from_bounds(min(self.bottom, other.bottom), max(self.top, other.top),
                               min(self.left, other.left), max(self.right, other.right))

    def __and__(self, other):
        """Overlap of the two, empty when there is none (bbox.py:126-149)."""
        bottom, top = max(self.bottom, other.bottom), min(self.top, other.top)
        left, right = max(self.left, other.left), min(self.right, other.right)
        if top < bottom or right < left:
            return Box((0, 0), width=0, height=0)
        return Box.from_bounds(bottom, top, left, right)

    def __eq__(self, other):
        return (self.left, self.right, self.top, self.bottom) == (other.left, other.right, other.top, other.bottom)

    def __ne__(self, other):
        return not self.__eq__(other)

    __hash__ = None

    def __str__(self):
        return "(({0}, {1}), ({2}, {3}))".format(self.bottom, self.top, self.left, self.right)

    def __repr__(self):
        return "<Box yx0={0}, height={1}, width={2}>".format(self.yx0, self.height, self.width)

    def copy(self):
        return Box((self.yx0[0], self.yx0[1]), self.height, self.width)


def trim_many(X, min_value=0):
    """Boxes of the pixels above `min_value` of every image of X [n, Ny, Nx] in one device pass; None where an image
    has no such pixel."""
    bounds, _, _ = _native.trim(np.asarray(X), float(min_value))
    return [None if b[0] < 0 else Box.from_bounds(int(b[0]), int(b[1]), int(b[2]), int(b[3])) for b in bounds]


def trim(X, min_value=0):
    """Bounding box of the pixels of the 2-D array X above `min_value` (bbox.py:170-191)."""
    box = trim_many(np.asarray(X)[None], min_value)[0]
    if box is None:
        # numpy's message for .min() of the empty selection the reference takes
        raise ValueError("zero-size array to reduction operation minimum which has no identity")
    return box


def flux_at_edge(X, min_value=0):
    """Whether a border pixel of X is above `min_value` (bbox.py:194-210)."""
    _, edge, _ = _native.trim(np.asarray(X)[None], float(min_value))
    return bool(edge[0])
