"""BlendBatch: many independent blends of one frame shape as flat arrays.

The reference has no batch notion (`Blend.fit` handles one scene, blend.py:65-102); this is the additive entry
point for BASELINE configs 3-4: the scene catalogue goes in as arrays, every source is initialised exactly as
`ExtendedSource.__init__` would (source.py:139-180,444-492) with the array operators batched on the device,
and `fit` runs `Blend.fit` for all blends in one launch sequence through the C ABI.
"""
import numpy as np

from . import _native, fft
from .source import build_detection_coadd, init_extended_morphs

_FLAGS_EXTENDED = (_native.F_CONSTRAINED | _native.F_KSPACE | _native.F_SYMMETRIC | _native.F_MONOTONIC)


class _Images:
    """What build_detection_coadd reads from an Observation."""
    def __init__(self, images):
        self.images = images


class BlendBatch:
    """B blends sharing (C, Ny, Nx), one set of band PSFs and one model PSF.

    images  float32 [B, C, Ny, Nx]; weights same shape or None (= 1, observation.py:148-151)
    psfs    [C, ky, kx] observed PSFs; model_psf [1, ky, kx] (Frame.psfs, observation.py:29-54)
    centers list of B integer arrays [K_b, 2] = (y, x): one ExtendedSource per row
    bg_rms  [C]
    Sources whose detection coadd has no pixel above the noise cut raise SourceInitError in the reference
    (source.py:171-173); here they are dropped and counted in `n_failed`; blends left without a source are
    dropped too (`kept` lists the surviving blend indices).
    """

    def __init__(self, images, psfs, model_psf, centers, bg_rms, weights=None, thresh=1.0,
                 symmetric=True, monotonic=True, stream=None):
        _native.require_device()
        images = np.ascontiguousarray(images, dtype=np.float32)
        B, C, Ny, Nx = images.shape
        self.shape = (C, Ny, Nx)
        bg_rms = np.asarray(bg_rms, dtype=np.float32)
        obs_psfs = fft.Fourier(np.array(psfs, dtype=np.float32))
        mod_psfs = fft.Fourier(np.array(model_psf, dtype=np.float32))
        obs_psfs.normalize(axes=(1, 2)) if not np.allclose(obs_psfs.sum(axis=(1, 2)), 1) else None
        mod_psfs.normalize(axes=(1, 2)) if not np.allclose(mod_psfs.sum(axis=(1, 2)), 1) else None
        diff = fft.match_psfs(obs_psfs, mod_psfs)                       # Observation.match, observation.py:191-194
        self.diff_kernels = diff.image
        spec = _native.kernel_spectrum(self.diff_kernels, Ny, Nx)
        peak_obs = obs_psfs.max(axis=(1, 2))
        peak_mod = mod_psfs[0].max()

        # ---- per-source SED + detection coadd (source.py:41-71,101-136), host
        seds, detects, cutoffs, ctrs, owner = [], [], [], [], []
        for b in range(B):
            ob = _Images(images[b])
            for (y, x) in np.asarray(centers[b]).reshape(-1, 2):
                sed = images[b, :, int(y), int(x)].copy()
                sed /= peak_obs
                sed = sed * peak_mod
                d, cut = build_detection_coadd(sed, bg_rms, ob, thresh)
                seds.append(sed); detects.append(d); cutoffs.append(cut); ctrs.append((int(y), int(x))); owner.append(b)
        # ---- sdss symmetry, monotonic(thresh=.1) on the device, noise cut, peak normalisation (source.py:160-179)
        morphs = np.empty((len(seds), Ny, Nx), np.float32)
        ok = np.zeros(len(seds), bool)
        step = 8192
        for s in range(0, len(seds), step):
            m, o = init_extended_morphs(np.array(detects[s:s + step]), cutoffs[s:s + step], ctrs[s:s + step], True, monotonic)
            morphs[s:s + step] = m
            ok[s:s + step] = o
        self.n_failed = int((~ok).sum())
        owner = np.array(owner)
        keep_src = np.nonzero(ok)[0]
        counts = np.bincount(owner[keep_src], minlength=B)
        self.kept = np.nonzero(counts > 0)[0]
        comp_off = np.concatenate([[0], np.cumsum(counts[self.kept])]).astype(np.int32)
        K = len(keep_src)
        flags = _FLAGS_EXTENDED
        if not symmetric:
            flags &= ~_native.F_SYMMETRIC
        if not monotonic:
            flags &= ~_native.F_MONOTONIC
        Bk = len(self.kept)
        self.host = dict(
            C=C, Ny=Ny, Nx=Nx, comp_off=comp_off,
            sed=np.array(seds, dtype=np.float32).reshape(-1, C)[keep_src],
            morph=morphs[keep_src],
            cflags=np.full(K, flags, np.uint32),
            center=np.array(ctrs, dtype=np.int32).reshape(-1, 2)[keep_src],
            shift=np.zeros((K, 2), np.float64),
            conv=np.full(K, 3, np.int32),
            images=images if Bk == B else images[self.kept],
            weights=None if weights is None else np.ascontiguousarray(weights, dtype=np.float32)[self.kept],
            kernel_spec=spec, ker_idx=np.tile(np.arange(C, dtype=np.int32), Bk),
            cw=np.asarray(mod_psfs[0].image, dtype=np.float32)[None], cw_idx=np.zeros(Bk, np.int32),
            it0=np.zeros(Bk, np.int32), prev_loss=np.zeros(Bk, np.float64))
        self.B, self.K = Bk, K
        self.mse = [[] for _ in range(Bk)]
        self._dev = _native.Batch(stream=stream, **self.host)
        # first constraint pass, it = 0 (source.py:492)
        self._dev.constrain()
        self._pull(self._dev.download())

    def _pull(self, out):
        h = self.host
        h["sed"], h["morph"], h["center"], h["shift"], h["conv"] = (out["sed"], out["morph"], out["center"],
                                                                   out["shift"], out["conv"])
        self.last = out

    @property
    def h2d_bytes(self):
        return self._dev.h2d_bytes

    @property
    def d2h_bytes(self):
        return self._dev.d2h_bytes

    def upload(self):
        """Host state -> device (what a fresh `fit` on host-resident blends costs)."""
        self._dev.upload(**{k: v for k, v in self.host.items() if k not in ("C", "Ny", "Nx")})

    def fit_device(self, max_iter=200, e_rel=1e-2, approximate_L=False, check_every=8, profile=False):
        """The loop only, on the state already in HBM; returns the timing dict of _native.Batch.fit."""
        return self._dev.fit(max_iter, e_rel, approximate_L, check_every, profile)

    def download(self, morph=True, out=None):
        return self._dev.download(morph=morph, out=out)

    def snapshot(self):
        """Keep a device-side copy of the current model state."""
        self._dev.snapshot()

    def restore(self):
        """Bring the snapshot back (stream-ordered device-to-device copy)."""
        self._dev.restore()

    def fit(self, max_iter=200, e_rel=1e-2, approximate_L=False, check_every=8):
        """`Blend.fit` for every blend: upload host state, run, read the results back into `self.host`;
        appends the per-iteration losses to `self.mse[b]`."""
        self.upload()
        info = self.fit_device(max_iter, e_rel, approximate_L, check_every)
        out = self.download()
        for b in range(self.B):
            n = int(out["n_iter"][b])
            self.mse[b].extend(float(v) for v in out["logL"][b, :n])
        self.host["it0"] = self.host["it0"] + out["n_iter"]
        self.host["prev_loss"] = np.array([m[-1] if m else 0. for m in self.mse])
        self._pull(out)
        return info

    def get_model(self, b):
        """Scene model of blend b, [C, Ny, Nx] (ComponentTree.get_model, component.py:321-347)."""
        k0, k1 = self.host["comp_off"][b], self.host["comp_off"][b + 1]
        return np.einsum("kc,kyx->cyx", self.host["sed"][k0:k1], self.host["morph"][k0:k1].reshape(k1 - k0, *self.shape[1:]))

    def close(self):
        self._dev.close()
