"""Pixel correspondence between a high- and a low-resolution frame (mirrors scarlet/resampling.py:3-125).
Setup work of LowResObservation.match, once per geometry, on the host."""
import numpy as np


class SourceInitError(Exception):
    """Datasets do not match, no overlap found (resampling.py:97-104)."""


def _pix2world(wcs, x, y):
    return wcs.all_pix2world(x, y, 0, ra_dec_order=True)[:2]


def _world2pix(wcs, ra, dec):
    return wcs.all_world2pix(ra, dec, 0, ra_dec_order=True)[:2]


def match_patches(shape_hr, shape_lr, wcs_hr, wcs_lr, isrot=True, perimeter='overlap', psf=False):
    """Coordinates of the low-resolution pixels that fall on the high-resolution frame, in both frames, and the
    high-resolution pixels inside the low-resolution frame (needed for PSF matching).

    Unrotated frames (isrot=False) are described per axis: rows and columns separately, like the reference, whose
    masks multiply the row and column conditions elementwise (resampling.py:90-92) -- both frames must be square.
    """
    assert perimeter in ['overlap', 'union'], 'perimeter should be either overlap or union.'
    assert wcs_hr is not None and wcs_lr is not None
    if np.size(shape_hr) not in (2, 3):
        raise ValueError('Wrong dimensions for reference image')
    if np.size(shape_lr) not in (2, 3):
        raise ValueError('Wrong dimensions for low resolution image')
    Ny_hr, Nx_hr = shape_hr[-2:]
    Ny_lr, Nx_lr = shape_lr[-2:]
    if isrot:
        # rotated frames: every low-resolution pixel is a point of its own (resampling.py:53-60), listed row-major
        rows_lr, cols_lr = [g.ravel() for g in np.indices((Ny_lr, Nx_lr))]
        rows_hr, cols_hr = np.arange(Ny_hr), np.arange(Nx_hr)
        ra_lr, dec_lr = _pix2world(wcs_lr, cols_lr, rows_lr)
        ra_hr, dec_hr = _pix2world(wcs_hr, cols_hr, rows_hr)
        cols_lr_in_hr, rows_lr_in_hr = _world2pix(wcs_hr, ra_lr, dec_lr)
        cols_hr_in_lr, rows_hr_in_lr = _world2pix(wcs_lr, ra_hr, dec_hr)
        inside_lr = ((cols_lr_in_hr >= 0) & (cols_lr_in_hr < Nx_hr + 1) & (rows_lr_in_hr >= 0) & (rows_lr_in_hr < Ny_hr + 1))
        inside_hr = ((cols_hr_in_lr >= 0) & (cols_hr_in_lr < Nx_lr + 1) & (rows_hr_in_lr >= 0) & (rows_hr_in_lr < Ny_lr + 1))
        if not inside_lr.any():
            raise SourceInitError
        coordhr_hr = (rows_hr[inside_hr], cols_hr[inside_hr])
        if perimeter == 'overlap':
            return ((rows_lr[inside_lr], cols_lr[inside_lr]),
                    (rows_lr_in_hr[inside_lr], cols_lr_in_hr[inside_lr]), coordhr_hr)
        return (rows_lr, cols_lr), (rows_lr_in_hr, cols_lr_in_hr), coordhr_hr
    if Ny_hr != Nx_hr or Ny_lr != Nx_lr:
        raise NotImplementedError("unrotated multi-resolution matching needs square frames (resampling.py:90-92)")
    rows_hr, cols_hr = np.arange(Ny_hr), np.arange(Nx_hr)
    rows_lr, cols_lr = np.arange(Ny_lr), np.arange(Nx_lr)
    ra_lr, dec_lr = _pix2world(wcs_lr, cols_lr, rows_lr)
    ra_hr, dec_hr = _pix2world(wcs_hr, cols_hr, rows_hr)
    cols_lr_in_hr, rows_lr_in_hr = _world2pix(wcs_hr, ra_lr, dec_lr)
    cols_hr_in_lr, rows_hr_in_lr = _world2pix(wcs_lr, ra_hr, dec_hr)
    inside_lr = ((cols_lr_in_hr >= 0) & (cols_lr_in_hr < Nx_hr + 1) & (rows_lr_in_hr >= 0) & (rows_lr_in_hr < Ny_hr + 1))
    inside_hr = ((cols_hr_in_lr >= 0) & (cols_hr_in_lr < Nx_lr + 1) & (rows_hr_in_lr >= 0) & (rows_hr_in_lr < Ny_lr + 1))
    coordhr_hr = (rows_hr[inside_hr], cols_hr[inside_hr])
    if not inside_lr.any():
        raise SourceInitError
    if perimeter == 'overlap':
        return ((rows_lr[inside_lr], cols_lr[inside_lr]),
                (rows_lr_in_hr[inside_lr], cols_lr_in_hr[inside_lr]), coordhr_hr)
    return (rows_lr, cols_lr), (rows_lr_in_hr, cols_lr_in_hr), coordhr_hr
