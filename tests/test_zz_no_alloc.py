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


def test_device_pointer_operators_match_host_pointer_ones(nat):
    """scl_*_dev (SURVEY 8(b) "Ownership": caller-owned device buffers, no hidden copies): monotonic sweep (bit-exact),
    max_pixel / centroid (exact centres) and the single-tile convolution on torch device tensors against the
    host-pointer entry points."""
    import ctypes
    import torch
    dev = torch.device("cuda:0")
    L = nat.lib()
    rng = np.random.default_rng(31)
    n, Ny, Nx = 6, 64, 64
    x = rng.random((n, Ny, Nx)).astype(np.float32)
    centers = np.stack([rng.integers(3, 61, n), rng.integers(3, 61, n)], axis=1).astype(np.int32)
    want = x.copy()
    nat.prox_weighted_monotonic(want, centers, 0.1)
    dx, dc = torch.from_numpy(x.copy()).to(dev), torch.from_numpy(centers).to(dev)
    p = lambda t: ctypes.c_void_p(t.data_ptr())          # noqa: E731
    nat._check(L.scl_prox_weighted_monotonic_f32_dev(p(dx), n, Ny, Nx, p(dc), ctypes.c_float(0.1), None))
    torch.cuda.synchronize()
    assert_array_equal(dx.cpu().numpy(), want)
    x64 = x.astype(np.float64)
    want64 = x64.copy()
    nat.prox_weighted_monotonic(want64, centers, 0.0)
    dx64 = torch.from_numpy(x64).to(dev)
    nat._check(L.scl_prox_weighted_monotonic_f64_dev(p(dx64), n, Ny, Nx, p(dc), ctypes.c_double(0.0), None))
    torch.cuda.synchronize()
    assert_array_equal(dx64.cpu().numpy(), want64)
    # measure
    psf = np.exp(-((np.arange(15) - 7)[:, None] ** 2 + (np.arange(15) - 7)[None, :] ** 2) / 4.0).astype(np.float32)
    c_want, s_want = nat.measure(x, centers, psf, True, True)
    dc2, ds = torch.from_numpy(centers.copy()).to(dev), torch.zeros((n, 2), dtype=torch.float64, device=dev)
    dxm, dpsf = torch.from_numpy(x).to(dev), torch.from_numpy(psf).to(dev)
    nat._check(L.scl_measure_f32_dev(p(dxm), n, Ny, Nx, p(dc2), p(ds), p(dpsf), 15, 15, 1, 1, None))
    torch.cuda.synchronize()
    assert_array_equal(dc2.cpu().numpy(), c_want)
    assert_allclose(ds.cpu().numpy(), s_want, rtol=0, atol=1e-12)
    # convolution
    ker = rng.random((2, 15, 15)).astype(np.float32)
    spec = nat.kernel_spectrum(ker, Ny, Nx)
    model = rng.random((3, 2, Ny, Nx)).astype(np.float32)
    idx = np.tile(np.arange(2, dtype=np.int32), 3)
    w = nat.convolve(model, spec, idx)
    dm, dout = torch.from_numpy(model).to(dev), torch.zeros((3, 2, Ny, Nx), dtype=torch.float32, device=dev)
    dspec = torch.from_numpy(np.ascontiguousarray(spec).view(np.float32)).to(dev)
    didx = torch.from_numpy(idx).to(dev)
    nat._check(L.scl_convolve_f32_dev(p(dm), p(dout), 3, 2, Ny, Nx, p(dspec), 2, p(didx), 0, None))
    torch.cuda.synchronize()
    assert_array_equal(dout.cpu().numpy(), w)
