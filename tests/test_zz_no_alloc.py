"""scl_prox_symmetry_noalloc_f32 (no counterpart in the reference; include/scarlet_b200.h): the symmetry operator on
caller-owned device arrays with its scratch in a static device array, against the existing host-pointer entry point
scl_prox_symmetry_f32 on the same inputs.  Collected last (file name) so that a failure here hides no other test."""
import numpy as np
import pytest
from numpy.testing import assert_allclose, assert_array_equal

pytestmark = pytest.mark.gpu

ATOL = 3e-6          # tests/test_gpu_ops.py::test_symmetry_random


@pytest.fixture(scope="module")
def nat():
    from scarlet_b200 import _native
    _native.require_device()
    return _native


def _case(seed, shape, n):
    rng = np.random.default_rng(seed)
    x = rng.random((n,) + shape).astype(np.float32)
    x[rng.random(x.shape) < .2] = 0
    centers = np.stack([rng.integers(3, shape[0] - 3, n), rng.integers(3, shape[1] - 3, n)], axis=1).astype(np.int32)
    shifts = rng.uniform(-.5, .5, (n, 2))
    return x, centers, shifts


def _noalloc(nat, torch, x, centers, shifts, mode, strength=1.0):
    dev = torch.device("cuda:0")
    dx = torch.from_numpy(x.copy()).to(dev)
    dc = torch.from_numpy(np.ascontiguousarray(centers, dtype=np.int32)).to(dev)
    ds = None if shifts is None else torch.from_numpy(np.ascontiguousarray(shifts, dtype=np.float64)).to(dev)
    bad = torch.zeros(1, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    n, Ny, Nx = x.shape
    nat.prox_symmetry_noalloc(dx.data_ptr(), n, Ny, Nx, dc.data_ptr(), None if ds is None else ds.data_ptr(), mode,
                              strength, bad.data_ptr(), None)        # legacy default stream
    torch.cuda.synchronize()
    return dx.cpu().numpy(), int(bad.item())


@pytest.mark.parametrize("shape,n", [((64, 64), 16), ((69, 57), 5), ((300, 280), 2)])
def test_noalloc_matches_existing_entry_point_twice(nat, shape, n):
    """Two calls in a row on different inputs (the second sees the first one's bytes in the shared scratch array)."""
    import torch
    for seed in (11, 12):
        x, centers, shifts = _case(seed, shape, n)
        want = x.copy()
        nat.prox_symmetry(want, centers, shifts, "kspace")
        got, bad = _noalloc(nat, torch, x, centers, shifts, "kspace")
        assert bad == 0
        assert_allclose(got, want, rtol=0, atol=ATOL)
        assert np.abs(got - x).max() > 1e-3           # the operator did something


def test_noalloc_flip_modes_whole_array_and_bad_centres(nat):
    import torch
    x, centers, shifts = _case(21, (40, 59), 6)
    for mode in ("soft", "sdss"):
        want = x.copy()
        nat.prox_symmetry(want, centers, None, mode, 1.0)
        got, bad = _noalloc(nat, torch, x, centers, None, mode)
        assert bad == 0
        assert_allclose(got, want, rtol=0, atol=ATOL)
    c2 = centers.copy()
    c2[0] = (-1, -1)                                   # whole array
    want = x.copy()
    nat.prox_symmetry(want, c2, shifts, "kspace")
    c2[3] = (40, 5)                                    # outside: the existing entry point refuses it on the host
    with pytest.raises(ValueError):
        nat.prox_symmetry(x.copy(), c2, shifts, "kspace")
    got, bad = _noalloc(nat, torch, x, c2, shifts, "kspace")
    assert bad == 1
    assert_array_equal(got[3], x[3])                   # left untouched, counted
    keep = [0, 1, 2, 4, 5]
    assert_allclose(got[keep], want[keep], rtol=0, atol=ATOL)


def test_noalloc_rejects_what_the_static_array_cannot_hold(nat):
    import torch
    dev = torch.device("cuda:0")
    dx = torch.zeros((480, 64, 64), dtype=torch.float32, device=dev)
    dc = torch.full((480, 2), 32, dtype=torch.int32, device=dev)
    ds = torch.zeros((480, 2), dtype=torch.float64, device=dev)
    with pytest.raises(NotImplementedError, match="SCL_NOALLOC_MAX_BYTES"):
        nat.prox_symmetry_noalloc(dx.data_ptr(), 480, 64, 64, dc.data_ptr(), ds.data_ptr())
    nat.prox_symmetry_noalloc(dx.data_ptr(), 479, 64, 64, dc.data_ptr(), ds.data_ptr())     # the stated maximum fits
    torch.cuda.synchronize()
    with pytest.raises(ValueError):
        nat.prox_symmetry_noalloc(dx.data_ptr(), 4, 64, 64, dc.data_ptr(), None, "kspace")
