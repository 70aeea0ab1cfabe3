"""BlendBatch: many independent blends of one frame shape as flat arrays.

The reference has no batch notion (`Blend.fit` handles one scene, blend.py:65-102); this is the additive entry
point for BASELINE configs 3-4: the scene catalogue goes in as arrays, every source is initialised exactly as
`ExtendedSource.__init__` would (source.py:139-180,444-492) on the device (`scl_batch_init_extended`: coadd, sdss
symmetry, monotonicity, noise cut, normalisation; only the [K, C] SED / weight arithmetic stays on the host),
and `fit` runs `Blend.fit` for all blends in one launch sequence through the C ABI.
"""
from concurrent.futures import ThreadPoolExecutor

import threading

import numpy as np

from . import _native, fft

_FLAGS_EXTENDED = (_native.F_CONSTRAINED | _native.F_KSPACE | _native.F_SYMMETRIC | _native.F_MONOTONIC)


_PER_COMP = ("sed", "morph", "cflags", "center", "shift", "conv")
_PER_BLEND = ("images", "weights", "cw_idx", "it0", "prev_loss")


class _Chunks:
    """The same interface as `_native.Batch`, over contiguous blend ranges that each own a `scl_batch` handle and a
    CUDA stream and are driven by one host thread each (the C calls release the GIL).  Blends are independent
    (SURVEY 8e), so the chunks never talk; running them side by side lets one chunk's kernels fill the SMs another
    leaves idle (tail waves, the thin per-blend kernels, the long tail of slowly converging blends) and overlaps
    one chunk's host<->device copies with another's iterations."""

    def __init__(self, n, host):
        self.C, self.Ny, self.Nx = host["C"], host["Ny"], host["Nx"]
        off = np.asarray(host["comp_off"])
        B = len(off) - 1
        n = max(1, min(int(n), B))
        # blend boundaries giving every chunk about the same number of components
        cuts = np.searchsorted(off, off[-1] * np.arange(1, n) / n)
        self.bounds = np.unique(np.concatenate([[0], cuts, [B]])).astype(int)
        self.koff = off[self.bounds].astype(int)
        self._psf_slices = {}
        self.streams = [_native.stream_create() for _ in range(len(self.bounds) - 1)]
        self.parts = [_native.Batch(stream=s, **self._slice(host, i)) for i, s in enumerate(self.streams)]
        for part in self.parts:
            part.wait_mode(True)            # several host threads per GPU (and several ranks per host): sleep, do not spin
        # the CUDA device is a per-thread setting: the worker threads adopt the device of the thread that built the batch
        self.pool = ThreadPoolExecutor(len(self.parts), initializer=_native.device_set, initargs=(_native.device_get(),))
        self.B, self.K = B, int(off[-1])
        self.max_iter = 0
        self.d2h_bytes = 0

    def _slice(self, a, i):
        """Chunk i's views of batch-wide arrays (dict keyed like the Batch constructor / download result)."""
        b0, b1 = self.bounds[i], self.bounds[i + 1]
        k0, k1 = self.koff[i], self.koff[i + 1]
        out = {}
        for key, v in a.items():
            if v is None or key in ("C", "Ny", "Nx", "kernel_spec", "cw", "ker_half"):
                out[key] = v
            elif key == "comp_off":
                out[key] = (np.asarray(v[b0:b1 + 1]) - k0).astype(np.int32)
            elif key == "ker_idx":
                out[key] = v[b0 * self.C:b1 * self.C]
                if a.get("psfs") is not None:                  # per-blend PSF sets: re-based to the chunk's own sets
                    out[key] = out[key] - b0 * self.C
            elif key == "psfs":             # one slice object per chunk, so that a re-upload keeps the matched spectra
                out[key] = self._psf_slices.setdefault((i, id(v)), dict(obs=v["obs"][b0:b1], model=v["model"]))
            elif key == "src_first":
                out[key] = (np.asarray(v[k0:k1]) - k0).astype(np.int32)
            elif key in _PER_COMP:
                out[key] = v[k0:k1]
            else:                                   # per blend: images, weights, cw_idx, it0, prev_loss, n_iter, ...
                out[key] = v[b0:b1]
        return out

    def _map(self, fn):
        return list(self.pool.map(fn, range(len(self.parts))))

    @property
    def h2d_bytes(self):
        return sum(p.h2d_bytes for p in self.parts)

    def upload(self, **host):
        self._map(lambda i: self.parts[i].upload(**{k: v for k, v in self._slice(host, i).items()
                                                    if k not in ("C", "Ny", "Nx", "ker_half")}))

    def fit(self, max_iter=200, e_rel=1e-2, approximate_L=False, check_every=8, profile=False, serial=False):
        """`serial=True` runs the chunks one after another (isolated per-kernel timings)."""
        run = (lambda i: self.parts[i].fit(max_iter, e_rel, approximate_L, check_every, profile))
        infos = [run(i) for i in range(len(self.parts))] if serial else self._map(run)
        self.max_iter = int(max_iter)
        agg = sum if serial else max
        return dict(device_ms=agg(r["device_ms"] for r in infos), steps=max(r["steps"] for r in infos),
                    kernel_ms=[sum(r["kernel_ms"][j] for r in infos) for j in range(4)],
                    launches=sum(r["launches"] for r in infos))

    def run(self, host, max_iter=200, e_rel=1e-2, approximate_L=False, check_every=8, out=None, morph=True):
        """upload -> fit -> download of every chunk on its own host thread and stream, with the PCIe link handed from
        chunk to chunk (one upload, one download at a time): a chunk iterates while the next one's inputs are still
        arriving, and its results leave while later chunks still iterate.  Returns (info, out) like fit + download."""
        if out is None:
            out = self.alloc_results(max_iter, morph)
        up, down = threading.Lock(), threading.Lock()

        def one(i):
            part = self.parts[i]
            with up:
                part.upload(**{k: v for k, v in self._slice(host, i).items() if k not in ("C", "Ny", "Nx", "ker_half")})
            r = part.fit(max_iter, e_rel, approximate_L, check_every, False)
            with down:
                part.download(morph=morph, out=self._slice(out, i))
            return r

        infos = self._map(one)
        self.max_iter = int(max_iter)
        self.d2h_bytes = sum(p.d2h_bytes for p in self.parts)
        return dict(device_ms=max(r["device_ms"] for r in infos), steps=max(r["steps"] for r in infos),
                    kernel_ms=[sum(r["kernel_ms"][j] for r in infos) for j in range(4)],
                    launches=sum(r["launches"] for r in infos)), out

    def constrain(self):
        self._map(lambda i: self.parts[i].constrain())

    def snapshot(self):
        self._map(lambda i: self.parts[i].snapshot())

    def restore(self):
        self._map(lambda i: self.parts[i].restore())

    def phase_clocks(self, enable=True):
        return [sum(v) for v in zip(*[p.phase_clocks(enable) for p in self.parts])]

    def products(self, min_value=0.0):
        parts = self._map(lambda i: self.parts[i].products(min_value))
        return tuple(np.concatenate([p[j] for p in parts]) for j in range(3))

    def render_components(self, k0=0, nk=None):
        nk = self.K - k0 if nk is None else int(nk)
        out = []
        for i, part in enumerate(self.parts):           # component ranges of the chunks overlapping [k0, k0 + nk)
            lo, hi = max(k0, self.koff[i]), min(k0 + nk, self.koff[i + 1])
            if hi > lo:
                out.append(part.render_components(lo - self.koff[i], hi - lo))
        return np.concatenate(out) if out else np.empty((0, self.C, self.Ny, self.Nx), np.float32)

    alloc_results = _native.Batch.alloc_results

    def download(self, morph=True, out=None):
        if out is None:
            out = self.alloc_results(self.max_iter, morph)
        self._map(lambda i: self.parts[i].download(morph=morph, out=self._slice(out, i)))
        self.d2h_bytes = sum(p.d2h_bytes for p in self.parts)
        return out

    def close(self):
        for p in self.parts:
            p.close()
        for s in self.streams:
            _native.stream_destroy(s)
        self.parts, self.streams = [], []
        self.pool.shutdown()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class BlendBatch:
    """B blends sharing (C, Ny, Nx), one set of band PSFs and one model PSF.

    images  float32 [B, C, Ny, Nx]; weights same shape or None (= 1, observation.py:148-151)
    psfs    [C, ky, kx] observed PSFs; model_psf [1, ky, kx] (Frame.psfs, observation.py:29-54)
    centers list of B integer arrays [K_b, 2] = (y, x): one ExtendedSource per row
    bg_rms  [C]
    Sources whose detection coadd has no pixel above the noise cut raise SourceInitError in the reference
    (source.py:171-173); here they are dropped and counted in `n_failed`; blends left without a source are
    dropped too (`kept` lists the surviving blend indices).
    streams > 1 splits the batch into that many independently driven chunks, one CUDA stream each (`_Chunks`).
    """

    def __init__(self, images, psfs, model_psf, centers, bg_rms, weights=None, thresh=1.0,
                 symmetric=True, monotonic=True, stream=None, streams=1, kinds=None, flux_percentiles=(25,),
                 detect_unit_weights=False):
        """kinds: list of B integer arrays [K_b] aligned with `centers` -- 0 ExtendedSource (default), 1
        MultiComponentSource (len(flux_percentiles) + 1 components, source.py:242-291,537-641), 2 PointSource
        (source.py:323-440); all of them initialised on the device (scl_batch_init_sources)."""
        _native.require_device()
        images = np.ascontiguousarray(images, dtype=np.float32)
        B, C, Ny, Nx = images.shape
        self.shape = (C, Ny, Nx)
        bg_rms = np.asarray(bg_rms, dtype=np.float32)
        psf_arr = np.array(psfs, dtype=np.float32)
        per_blend = psf_arr.ndim == 4                   # [B, C, ky, kx]: every blend has its own observed PSFs
        mod_psfs = fft.Fourier(np.array(model_psf, dtype=np.float32))
        mod_psfs.normalize(axes=(1, 2)) if not np.allclose(mod_psfs.sum(axis=(1, 2)), 1) else None
        if per_blend:
            if psf_arr.shape[:2] != (B, C):
                raise ValueError("per-blend psfs must be [B, C, ky, kx]")
            tot = psf_arr.sum(axis=(2, 3))
            if not np.allclose(tot, 1):
                psf_arr = psf_arr * (1 / tot)[:, :, None, None]          # Frame.__init__ normalises, observation.py:48-52
            # Observation.match runs inside the batch (scl_batch_desc.psf_obs): the B * C spectra never leave HBM
            self.diff_kernels, spec = None, None
            peak_obs = psf_arr.max(axis=(2, 3))                          # [B, C]
        else:
            obs_psfs = fft.Fourier(psf_arr)
            obs_psfs.normalize(axes=(1, 2)) if not np.allclose(obs_psfs.sum(axis=(1, 2)), 1) else None
            diff, spec = _native.match_psfs(obs_psfs.image[None], mod_psfs.image[:1], frame_shape=(Ny, Nx))
            self.diff_kernels, spec = diff[0], spec[0]                  # Observation.match, observation.py:191-194
            peak_obs = obs_psfs.max(axis=(1, 2))
        peak_mod = mod_psfs[0].max()

        # ---- per-source SED and coadd weights (source.py:41-71,101-136): [K, C] arrays, vectorised over all
        # sources in the dtypes the reference's scalar code produces for float32 images / bg_rms
        counts0 = np.array([len(np.asarray(c).reshape(-1, 2)) for c in centers], dtype=np.int64)
        owner = np.repeat(np.arange(B), counts0)
        ctr = (np.concatenate([np.asarray(c).reshape(-1, 2) for c in centers]).astype(np.int32)
               if counts0.sum() else np.zeros((0, 2), np.int32))
        kind_src = (np.concatenate([np.asarray(k, dtype=np.int32).reshape(-1) for k in kinds]) if kinds is not None
                    else np.zeros(len(ctr), np.int32))
        if len(kind_src) != len(ctr) or np.any((kind_src < 0) | (kind_src > 2)):
            raise ValueError("kinds: one entry of 0 (extended), 1 (multi-component) or 2 (point) per catalogue row")
        perc = np.sort(np.asarray(flux_percentiles, dtype=np.float32))
        n_layers = len(perc) + 1
        # catalogue rows -> components: a multi-component source owns n_layers consecutive components
        reps = np.where(kind_src == 1, n_layers, 1)
        first = np.concatenate([[0], np.cumsum(reps)])[:-1]
        owner, ctr, kind = np.repeat(owner, reps), np.repeat(ctr, reps, axis=0), np.repeat(kind_src, reps)
        src_first = np.repeat(first, reps).astype(np.int32)
        sed = images[owner, :, ctr[:, 0], ctr[:, 1]].astype(np.float32)
        sed /= np.asarray(peak_obs[owner] if per_blend else peak_obs, dtype=np.float32)
        # get_psf_sed scales with the model PSF's peak (source.py:64-69); PointSource.__init__ does not (source.py:377-381)
        sed = np.where((kind == 2)[:, None], sed, sed * np.float32(peak_mod)).astype(np.float32)
        if np.any(bg_rms <= 0):
            raise ValueError("bg_rms must be greater than zero in all channels")
        pos = sed > 0
        bg2 = bg_rms ** 2
        cwt = np.where(pos, sed / bg2, np.float32(0)).astype(np.float32)
        jterm = np.where(pos, sed ** 2 / bg2, np.float32(0)).astype(np.float32)
        cterm = (cwt ** 2 * bg2).astype(np.float32)
        jac = np.zeros(len(sed), np.float32)
        csum = np.zeros(len(sed), np.float32)
        for c in range(C):                                  # band-order float32 sums, like np.array([...]).sum()
            jac = jac + jterm[:, c]
            csum = csum + cterm[:, c]
        with np.errstate(divide="ignore", invalid="ignore"):
            cutoff = (np.float32(thresh) * np.sqrt(csum) / jac).astype(np.float32)
        # ---- morphologies on the device (source.py:139-180): coadd, sdss symmetry, monotonic(thresh=.1), noise cut,
        # peak normalisation -- one CTA per candidate source on the batch's own images (scl_batch_init_extended)
        Kc = len(sed)
        flags = np.where(kind == 1, _native.F_CONSTRAINED | _native.F_SHARED_CENTER | _native.F_SYMMETRIC | _native.F_MONOTONIC,
                         _FLAGS_EXTENDED).astype(np.uint32)
        if not symmetric:
            flags &= ~np.uint32(_native.F_SYMMETRIC)
        if not monotonic:
            flags &= ~np.uint32(_native.F_MONOTONIC)
        self.kind = kind
        wts = None if weights is None else np.ascontiguousarray(weights, dtype=np.float32)
        # A weight array that is 1 everywhere is the reference's `weights=None` (observation.py:148-151: W = 1), bit for
        # bit: with detect_unit_weights=True checked once here, and then neither stored on the device nor uploaded
        # (opt-in: the bench keeps the per-pixel array its workload names)
        self.uniform_weights = bool(detect_unit_weights and wts is not None and wts.size > 0 and
                                    float(wts.min()) == 1.0 and float(wts.max()) == 1.0)
        if self.uniform_weights:
            wts = None

        def host_dict(keep_src, kept, morph, sed=sed):
            cnt = np.bincount(owner[keep_src], minlength=B)[kept]
            # first component of each kept component's source, renumbered after dropping
            newpos = -np.ones(Kc + 1, np.int64)
            newpos[keep_src] = np.arange(len(keep_src))
            return dict(
                C=C, Ny=Ny, Nx=Nx, comp_off=np.concatenate([[0], np.cumsum(cnt)]).astype(np.int32),
                sed=sed[keep_src], morph=morph, cflags=flags[keep_src],
                **(dict(src_first=newpos[src_first[keep_src]].astype(np.int32)) if np.any(kind == 1) else {}),
                center=ctr[keep_src], shift=np.zeros((len(keep_src), 2), np.float64),
                conv=np.full(len(keep_src), 3, np.int32),
                images=images if len(kept) == B else images[kept],
                weights=None if wts is None else (wts if len(kept) == B else wts[kept]),
                kernel_spec=spec,
                ker_idx=(np.arange(C * len(kept), dtype=np.int32) if per_blend
                         else np.tile(np.arange(C, dtype=np.int32), len(kept))),
                **(dict(psfs=dict(obs=psf_arr if len(kept) == B else psf_arr[kept],
                                  model=np.asarray(mod_psfs.image[:1], dtype=np.float32))) if per_blend else {}),
                # reach of the difference kernels: frames beyond one convolution tile run in scene mode
                ker_half=_native.kernel_half(psf_arr.shape[-2], psf_arr.shape[-1], Ny, Nx),
                cw=np.asarray(mod_psfs[0].image, dtype=np.float32)[None], cw_idx=np.zeros(len(kept), np.int32),
                it0=np.zeros(len(kept), np.int32), prev_loss=np.zeros(len(kept), np.float64))

        has = np.nonzero(counts0 > 0)[0]
        cand = host_dict(np.arange(Kc), has, None)
        dev = _native.Batch(stream=stream, **cand)
        ok = (dev.init_sources(kind if kinds is not None else None, cwt, jac, cutoff, True, monotonic, perc)
              if Kc else np.zeros(0, bool))
        self.n_failed = int((~ok).sum())
        keep_src = np.nonzero(ok)[0]
        self.kept = np.nonzero(np.bincount(owner[keep_src], minlength=B) > 0)[0]
        if self.n_failed == 0 and streams <= 1:
            self.host = cand
            self._dev = dev
        else:
            # sources without a pixel above the noise cut are dropped (source.py:171-173), and with them blends left
            # without a source: rebuild the batch from the surviving candidates
            first_out = dev.download(morph=True)
            dev.close()
            self.kind = kind[keep_src]
            self.host = host_dict(keep_src, self.kept, first_out["morph"][keep_src], first_out["sed"])
            self._dev = None
        Bk, K = len(self.kept), len(keep_src)
        self.B, self.K = Bk, K
        self.mse = [[] for _ in range(Bk)]
        if self._dev is None:
            self._dev = _Chunks(streams, self.host) if streams > 1 else _native.Batch(stream=stream, **self.host)
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
        self._dev.upload(**{k: v for k, v in self.host.items() if k not in ("C", "Ny", "Nx", "ker_half")})

    def fit_device(self, max_iter=200, e_rel=1e-2, approximate_L=False, check_every=8, profile=False, serial=False):
        """The loop only, on the state already in HBM; returns the timing dict of _native.Batch.fit.
        `serial` (chunked batches only): one chunk after another instead of side by side."""
        if isinstance(self._dev, _Chunks):
            return self._dev.fit(max_iter, e_rel, approximate_L, check_every, profile, serial)
        return self._dev.fit(max_iter, e_rel, approximate_L, check_every, profile)

    def download(self, morph=True, out=None):
        return self._dev.download(morph=morph, out=out)

    def snapshot(self):
        """Keep a device-side copy of the current model state."""
        self._dev.snapshot()

    def restore(self):
        """Bring the snapshot back (stream-ordered device-to-device copy)."""
        self._dev.restore()

    def fit_host(self, max_iter=200, e_rel=1e-2, approximate_L=False, check_every=8, out=None, morph=True):
        """Host state -> device, the loop, results -> host arrays (`out` as from `alloc_results`, e.g. page-locked);
        a chunked batch (`streams > 1`) pipelines the three steps chunk by chunk (`_Chunks.run`).  Returns (info, out);
        `self.host` is left untouched (see `fit`)."""
        if isinstance(self._dev, _Chunks):
            return self._dev.run(self.host, max_iter, e_rel, approximate_L, check_every, out, morph)
        self.upload()
        info = self.fit_device(max_iter, e_rel, approximate_L, check_every)
        return info, self.download(morph=morph, out=out)

    def fit(self, max_iter=200, e_rel=1e-2, approximate_L=False, check_every=8):
        """`Blend.fit` for every blend: upload host state, run, read the results back into `self.host`;
        appends the per-iteration losses to `self.mse[b]`."""
        info, out = self.fit_host(max_iter, e_rel, approximate_L, check_every)
        for b in range(self.B):
            n = int(out["n_iter"][b])
            self.mse[b].extend(float(v) for v in out["logL"][b, :n])
        self.host["it0"] = self.host["it0"] + out["n_iter"]
        self.host["prev_loss"] = np.array([m[-1] if m else 0. for m in self.mse])
        self._pull(out)
        return info

    def products(self, min_value=0.0):
        """What a caller reads out of the fitted blends, per component and from the device state:
        flux [K, C] (Component.get_flux, component.py:172-175), bounds [K, 4] = (bottom, top, left, right) of the
        morphology above `min_value` (bbox.trim) or -1, flags [K] (BlendFlag bits: convergence as left by the fit,
        EDGE_PIXELS, NO_VALID_PIXELS)."""
        return self._dev.products(min_value)

    def render_components(self, k0=0, nk=None):
        """Observation.render(component.get_model()) for components [k0, k0 + nk), float32 [nk, C, Ny, Nx]: the
        per-source templates in the observed frame (observation.py:198-220)."""
        return self._dev.render_components(k0, nk)

    def get_model(self, b):
        """Scene model of blend b, [C, Ny, Nx] (ComponentTree.get_model, component.py:321-347)."""
        k0, k1 = self.host["comp_off"][b], self.host["comp_off"][b + 1]
        return np.einsum("kc,kyx->cyx", self.host["sed"][k0:k1], self.host["morph"][k0:k1].reshape(k1 - k0, *self.shape[1:]))

    def close(self):
        self._dev.close()
