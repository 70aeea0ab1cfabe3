"""Stand-ins that let the UNMODIFIED reference package at /root/reference be
imported in the build container (which lacks autograd, proxmin, Eigen and
astropy and ships numpy 2 / scipy 1.18).  Used only by make_golden.py to
generate the fixtures in this directory; never imported by tests or product.

What is substituted:
  * numpy/scipy compat: `np.int` alias; float arguments to
    scipy.fftpack.helper.next_fast_len.
  * proxmin.operators.prox_plus/prox_hard/prox_soft: restated from their
    published definitions (pinned by reference tests/test_update.py:23-97).
  * scarlet.operators_pybind11: pure-Python loops with the loop bodies of
    operators_pybind11.cc:11-50 (typed accumulation in the array dtype).
  * autograd.grad: returns the analytic adjoint of Blend._loss evaluated with
    the reference's own fft module (forward model, loss and every other step
    are the reference's code).
"""
import sys
import types
import numpy as np


def install(reference_root="/root/reference"):
    if not hasattr(np, "int"):
        np.int = int
    import warnings
    warnings.filterwarnings("ignore")
    from scipy import fftpack
    import scipy.fftpack.helper as helper
    _orig = helper.next_fast_len
    helper.next_fast_len = lambda n: _orig(int(n))
    fftpack.helper = helper

    # ---- autograd
    ag = types.ModuleType("autograd")
    agnp = types.ModuleType("autograd.numpy")
    agnp.__dict__.update(np.__dict__)
    boxes = types.ModuleType("autograd.numpy.numpy_boxes")

    class ArrayBox:  # never instantiated
        pass
    boxes.ArrayBox = ArrayBox
    agnp.numpy_boxes = boxes

    def grad(fun, argnums):
        blend = fun.__self__

        def g(*params):
            import scarlet.fft as rfft
            K = blend.K
            seds, morphs = params[:K], params[K:]
            model = blend.get_model(seds, morphs)
            total = 0
            G = np.zeros(model.shape, dtype=model.dtype)
            for obs in blend.observations:
                total = total + obs.get_loss(model)
                m = obs.render(model)
                r = obs.weights * obs.weights * (m - obs.images)
                if obs._diff_kernels is not None:
                    r = rfft._kspace_operation(rfft.Fourier(r), obs._diff_kernels, 3,
                                               lambda a, b: a * np.conj(b), r.shape, axes=(1, 2)).image
                G[obs._band_slice] += r
            blend.mse.append(total)
            sg = [np.sum(G * morphs[k][None], axis=(1, 2)) for k in range(K)]
            mg = [np.sum(seds[k][:, None, None] * G, axis=0) for k in range(K)]
            return tuple(sg + mg)
        return g
    ag.grad = grad
    ag.numpy = agnp
    sys.modules.update({"autograd": ag, "autograd.numpy": agnp, "autograd.numpy.numpy_boxes": boxes})

    # ---- proxmin
    pm = types.ModuleType("proxmin")
    ops = types.ModuleType("proxmin.operators")
    utils = types.ModuleType("proxmin.utils")

    def prox_plus(X, step):
        X[X < 0] = 0
        return X

    def prox_hard(X, step, thresh=0):
        X[np.abs(X) < thresh * step] = 0
        return X

    def prox_soft(X, step, thresh=0):
        t = thresh * step
        X[:] = np.sign(X) * np.maximum(np.abs(X) - t, 0)
        return X

    def prox_unity_plus(X, step, axis=0):
        X[X < 0] = 0
        X[:] = X / np.sum(X, axis=axis, keepdims=True)
        return X
    ops.prox_plus, ops.prox_hard, ops.prox_soft, ops.prox_unity_plus = prox_plus, prox_hard, prox_soft, prox_unity_plus

    class MatrixAdapter:
        def __init__(self, L, axis=None):
            self.L, self.axis = L, axis

        @property
        def spectral_norm(self):
            return None
    utils.MatrixAdapter = MatrixAdapter
    pm.operators, pm.utils = ops, utils
    sys.modules.update({"proxmin": pm, "proxmin.operators": ops, "proxmin.utils": utils})

    # ---- native module
    nat = types.ModuleType("scarlet.operators_pybind11")

    def prox_monotonic(X, step, ref_idx, dist_idx, thresh):
        for d in dist_idx:
            X[d] = min(X[d], X[ref_idx[d]] * (1 - thresh))

    def prox_weighted_monotonic(flat, step, weights, offsets, dist_idx, thresh):
        T = flat.dtype.type
        w = np.asarray(weights, dtype=flat.dtype)
        one_m = T(T(1) - T(thresh))
        for d in dist_idx:
            ref = T(0)
            for i in range(len(offsets)):
                if w[i, d] > 0:
                    ref = T(ref + T(flat[offsets[i] + d] * w[i, d]))
            r = T(ref * one_m)
            if r < flat[d]:
                flat[d] = r
    nat.prox_monotonic, nat.prox_weighted_monotonic = prox_monotonic, prox_weighted_monotonic
    sys.modules["scarlet.operators_pybind11"] = nat

    sys.path.insert(0, reference_root)
    import scarlet
    scarlet.operators_pybind11 = nat
    return scarlet
