"""A minimal affine WCS exposing exactly the attributes scarlet's multi-resolution path touches
(observation.py:91-94,282-296,426-432; resampling.py:67-87): `naxis`, `array_shape`, `wcs.pc / crval / crpix`,
`deepcopy()`, `all_pix2world`, `all_world2pix`, `wcs_world2pix`.  astropy is absent in the build container
(SURVEY.md Appendix B); any object with this surface (astropy.wcs.WCS included) works with LowResObservation.

world = crval + pc @ (pixel - (crpix - 1))   for origin 0, pixel = (x, y), world = (ra, dec)
"""
import copy
import numpy as np


class _Inner:
    def __init__(self, pc, crval, crpix):
        self.pc = np.array(pc, dtype=np.float64)
        self.crval = tuple(crval)
        self.crpix = tuple(crpix)


class AffineWCS:
    naxis = 2

    def __init__(self, pc, crpix, crval=(0., 0.), array_shape=None):
        self.wcs = _Inner(pc, crval, crpix)
        self.array_shape = array_shape

    def deepcopy(self):
        return copy.deepcopy(self)

    def _split(self, args, kw):
        return np.asarray(args[0], dtype=np.float64), np.asarray(args[1], dtype=np.float64), args[2]

    def all_pix2world(self, *args, **kw):
        x, y, origin = self._split(args, kw)
        pc = self.wcs.pc
        cx, cy = np.asarray(self.wcs.crpix[:2], dtype=np.float64) - 1 + origin
        dx, dy = x - cx, y - cy
        if dx.shape != dy.shape:     # resampling.py:73-76 passes all columns and all rows of a non-square frame
            assert pc[0, 1] == 0 and pc[1, 0] == 0, "separate axes need an unrotated frame"
            return self.wcs.crval[0] + pc[0, 0] * dx, self.wcs.crval[1] + pc[1, 1] * dy
        return (self.wcs.crval[0] + pc[0, 0] * dx + pc[0, 1] * dy,
                self.wcs.crval[1] + pc[1, 0] * dx + pc[1, 1] * dy)

    def all_world2pix(self, *args, **kw):
        ra, dec, origin = self._split(args, kw)
        inv = np.linalg.inv(self.wcs.pc)
        cx, cy = np.asarray(self.wcs.crpix[:2], dtype=np.float64) - 1 + origin
        da, dd = ra - self.wcs.crval[0], dec - self.wcs.crval[1]
        if da.shape != dd.shape:
            assert inv[0, 1] == 0 and inv[1, 0] == 0, "separate axes need an unrotated frame"
            return cx + inv[0, 0] * da, cy + inv[1, 1] * dd
        return (cx + inv[0, 0] * da + inv[0, 1] * dd, cy + inv[1, 0] * da + inv[1, 1] * dd)

    def wcs_world2pix(self, a, b, origin, *rest):
        """Frame.get_pixel (observation.py:85-97) uses the two return values as (y, x) in the order the sky
        coordinate was given; with this stub a sky coordinate is (world_y, world_x) and the answer (pix_y, pix_x)."""
        x, y = self.all_world2pix(b, a, origin)
        return np.asarray(y), np.asarray(x)


def make_pair(hr_shape, lr_shape, scale=4.0, angle_deg=0.0, lr_offset=(0.3, -0.2)):
    """HR frame (pixel scale 1) and an LR frame (pixel scale `scale`, rotated by angle_deg) looking at the same
    patch: both reference pixels at their frame centres (+ a sub-pixel offset of the LR grid, in LR pixels)."""
    (hy, hx), (ly, lx) = hr_shape, lr_shape
    whr = AffineWCS(np.eye(2), crpix=(hx / 2 + 1, hy / 2 + 1), array_shape=(hy, hx))
    t = np.deg2rad(angle_deg)
    rot = np.array([[np.cos(t), -np.sin(t)], [np.sin(t), np.cos(t)]])
    wlr = AffineWCS(scale * rot, crpix=(lx / 2 + 1 + lr_offset[1], ly / 2 + 1 + lr_offset[0]), array_shape=(ly, lx))
    return whr, wlr
