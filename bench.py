#!/usr/bin/env python
"""bench.py -- blends/sec of `Blend.fit` on BASELINE config 3 (synthetic 5-band 64x64 blends, 2-8 ExtendedSources).

One "step" = Blend.fit(max_iter=200, e_rel=1e-2) (the reference's defaults, blend.py:65) over the whole batch of
blends resident on one GPU.  Ranks own disjoint blends (weak scaling: --blends per GPU), no collective inside the
step; per-blend result records are all-gathered over NCCL after the step.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--blends B] [--impl b200|reference]
  python bench.py --workload config4 [--steps K] [--warmup W]      (one 6-band 512x512 scene, ~330 components)

Prints ONE JSON line (rank 0).  `value` = blends/s with inputs resident in HBM; `e2e` = the same through the
C ABI with host buffers (upload + fit + download per step); `roofline` = the dominant kernel against the
measured HBM peak; `cpu_baseline` = the CPU oracle (restatement of the reference, oracle/) on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "blends_per_sec"
UNIT = "blends/s"
WORKLOAD = "config3: synthetic 5-band 64x64 blends, 2-8 ExtendedSources, 41x41 PSFs, Blend.fit(200, e_rel=1e-2)"
C, NY, NX, PSF = 5, 64, 64, 41
MAX_ITER, E_REL = 200, 1e-2


# ------------------------------------------------------------------------------------------ CPU side (oracle)
def _cpu_worker(args):
    """Fit a list of blends with the CPU oracle; returns (n_blends, n_iters, seconds in fit)."""
    os.environ["OMP_NUM_THREADS"] = "1"
    images, weights, psfs, mpsf, centers, max_iter, e_rel = args
    from oracle import ref_blend as rb
    tot_it, tot_t, nb = 0, 0.0, 0
    for b in range(len(images)):
        frame = rb.Frame(images[b].shape, psfs=mpsf, dtype=np.float32)
        obs = rb.Observation(images[b], psfs=psfs, weights=weights[b]).match(frame)
        bg = np.ones(images.shape[1], np.float32)
        srcs = []
        for (y, x) in centers[b]:
            try:
                srcs.append(rb.ExtendedSource(frame, (int(y), int(x)), obs, bg))
            except rb.SourceInitError:
                pass
        if not srcs:
            continue
        blend = rb.Blend(srcs, obs)
        t0 = time.perf_counter()
        blend.fit(max_iter, e_rel)
        tot_t += time.perf_counter() - t0
        tot_it += blend.it
        nb += 1
    return nb, tot_it, tot_t


def cpu_baseline(per_worker=24, max_workers=32, first_index=0):
    """Oracle Blend.fit on a bounded sample of the same workload, one process per host core."""
    import multiprocessing as mp
    from scarlet_b200 import synth
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])
    cores = min(max_workers, len(os.sched_getaffinity(0)))
    n = cores * per_worker
    data = synth.make_blends(range(first_index, first_index + n), C, NY, NX, psf_size=PSF, device="cpu")
    jobs = []
    for w in range(cores):
        sl = slice(w * per_worker, (w + 1) * per_worker)
        jobs.append((data["images"][sl], data["weights"][sl], data["psfs"], data["model_psf"], data["centers"][sl],
                     MAX_ITER, E_REL))
    ctx = mp.get_context("spawn")
    t0 = time.perf_counter()
    with ctx.Pool(cores) as pool:
        res = pool.map(_cpu_worker, jobs)
    wall = time.perf_counter() - t0
    nb = sum(r[0] for r in res)
    its = sum(r[1] for r in res)
    tmax = max(r[2] for r in res)
    return dict(value=nb / tmax, unit=UNIT, cores=cores, kind="port",
                sample="%d blends of the same generator (indices %d..%d), %d per process; %.1f s of fit time on the "
                       "slowest process (pool wall %.1f s incl. source init); %.1f iterations/blend"
                       % (nb, first_index, first_index + n - 1, per_worker, tmax, wall, its / max(nb, 1)),
                iters_per_sec=its / tmax, single_core_blends_per_sec=nb / sum(r[2] for r in res),
                balanced_value=nb * cores / sum(r[2] for r in res), fit_wall_s=tmax)


# ------------------------------------------------------------------------------------------ ncu evidence
KERNEL_SOURCES = {
    "scl_k_forward_adjoint": ["k_forward.cu", "fft_reg.cuh", "fft_smem.cuh", "common.cuh"],
    "scl_k_prox": ["k_prox.cu", "prox_dev.cuh", "k_update.cu", "common.cuh"],
    "scl_k_lipschitz": ["k_grad.cu", "common.cuh"],
}


def sources_sha(kernel):
    """sha256 (first 16 hex digits) of the source files a kernel is compiled from."""
    import hashlib
    h = hashlib.sha256()
    for f in KERNEL_SOURCES[kernel]:
        with open(os.path.join(ROOT, "scarlet_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def kernel_evidence():
    """profiles/kernel_evidence.json entries whose source hash matches the tree (see tools/ncu_stamp.py)."""
    try:
        ev = json.load(open(os.path.join(ROOT, "profiles", "kernel_evidence.json")))
    except (OSError, ValueError):
        return {}
    out = {}
    for name, e in ev.get("kernels", {}).items():
        if name in KERNEL_SOURCES and e.get("sources_sha") == sources_sha(name):
            out[name] = dict(traffic=e.get("traffic"), issue_frac=e.get("issue_frac"), fma_pipe_frac=e.get("fma_pipe_frac"),
                             stamp="%s, sources %s, commit %s" % (e.get("profile"), e.get("sources_sha"), e.get("commit")))
    return out


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi -lms 200 during the timed region (profiling recipe's clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([f.strip() for f in line.split(",")])

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.proc.terminate()
        self.th.join(timeout=2)
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    power_w_max=max(pw) if pw else None, samples=len(sm), reasons=sorted(reasons))


# ------------------------------------------------------------------------------------------ GPU arm
def alg_bytes(comp_off, n_iter=None, weight_planes=1):
    """Algorithmic bytes per launch of each kernel, summed over the blends of one batch.
    With n_iter: summed over the iterations each blend actually ran (launches differ in active blends)."""
    K = np.diff(comp_off).astype(np.float64)
    N = NY * NX
    w = np.ones_like(K) if n_iter is None else n_iter.astype(np.float64)
    Cw = C * weight_planes                                            # 0 when the weights are the scalar 1 (SURVEY 8(d) C_w)
    k1 = 4 * N * (K + 2 * C + Cw) + 4 * C * PSF * PSF + 8 * K * C     # S, images, weights in; G out; diff kernel; A, dA
    k2 = 12 * K * C + 8 * K * K + 8 * C + 16                          # A, dA in; A out; Gram, loss parts in; logL, L out
    k3 = 4 * N * (2 * K + C) + 8 * K * C                              # S in/out, G in; A in/out
    it = 4 * N * (2 * K + C + Cw) + 4 * C * PSF * PSF + 8 * K * C     # fused-iteration figure B_iter
    return [float((k1 * w).sum()), float((k2 * w).sum()), float((k3 * w).sum())], float((it * w).sum())


def bind_near_gpu(local):
    """Restrict this rank (and the chunk threads it spawns) to the host cores on the GPU's own NUMA node
    (/sys/bus/pci/devices/<bus id>/local_cpulist), so that the staging buffers allocated afterwards are first touched
    there and eight ranks do not all stage through node 0.  Returns a description for the bench line."""
    try:
        bus = subprocess.check_output(["nvidia-smi", "-i", str(local), "--query-gpu=pci.bus_id", "--format=csv,noheader"],
                                      text=True).strip().lower()
        bus = bus[-12:] if len(bus) > 12 else bus                      # sysfs uses a 4-digit PCI domain
        base = "/sys/bus/pci/devices/" + bus
        node = int(open(base + "/numa_node").read())
        cpus = set()
        for part in open(base + "/local_cpulist").read().strip().split(","):
            if part:
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
        have = os.sched_getaffinity(0)
        use = cpus & have
        if node < 0 or not use or use == have:
            return "gpu %d: numa node %d, local cpus %d, allowed %d: affinity left as is" % (local, node, len(cpus), len(have))
        os.sched_setaffinity(0, use)
        return "gpu %d: bound to %d cores of numa node %d" % (local, len(use), node)
    except (OSError, ValueError, subprocess.SubprocessError) as e:
        return "affinity left as is (%s)" % type(e).__name__


def run_b200(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    binding = bind_near_gpu(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm_peak, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)") if "hbm_gbs" in peaks \
        else (6650.0, "fallback (B200_PROFILING.md)")

    import __graft_entry__
    if rank == 0:
        __graft_entry__.build()
    if world > 1:
        dist.barrier()
    from scarlet_b200 import synth, shard
    from scarlet_b200.batch import BlendBatch

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline()

    # ---- inputs: rank r owns blend indices [r*B, (r+1)*B)
    B = args.blends
    t0 = time.perf_counter()
    data = synth.make_blends(shard.weak_range(B, rank), C, NY, NX, psf_size=PSF)
    t_gen = time.perf_counter() - t0
    t0 = time.perf_counter()
    psfs = data["psfs"]
    if args.per_blend_psfs:
        psfs = np.broadcast_to(np.asarray(psfs, np.float32), (len(data["images"]),) + np.shape(psfs)).copy()
    bb = BlendBatch(data["images"], psfs, data["model_psf"], data["centers"], data["bg_rms"],
                    weights=data["weights"], streams=args.streams)
    t_init = time.perf_counter() - t0
    bb.snapshot()
    comp_off = bb.host["comp_off"]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def allmax(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- device-resident throughput: restore the initial state, run the loop
    for _ in range(args.warmup):
        bb.restore()
        bb.fit_device(MAX_ITER, E_REL, check_every=args.check_every)
    sampler = ClockSampler(local) if rank == 0 else None
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kms = np.zeros(4)
    launches = steps_run = 0
    barrier()
    ev0.record()
    for _ in range(args.steps):
        bb.restore()
        info = bb.fit_device(MAX_ITER, E_REL, check_every=args.check_every, profile=True)
        kms += info["kernel_ms"]
        launches += info["launches"]
        steps_run += info["steps"]
    ev1.record()
    barrier()
    ms = allmax(ev0.elapsed_time(ev1))
    clocks = sampler.stop() if sampler else None
    out = bb.download(morph=False)
    n_iter = out["n_iter"]
    blends_total = allsum(float(bb.B))
    iters_total = allsum(float(n_iter.sum()))
    value = blends_total * args.steps / (ms / 1e3)

    # ---- end to end through the C ABI with page-locked host buffers: upload + fit + download (+ gather)
    # The public call (BlendBatch.fit_host) pipelines upload / loop / download chunk by chunk when the batch is split
    # into independently driven chunks; the device-resident number above is best with one handle, so the end-to-end
    # leg gets its own batch of the same blends with --e2e-streams chunks.
    cudart = torch.cuda.cudart()
    bbe = bb
    if args.e2e_streams != args.streams:
        # the public entry point recognises a weight array of ones as the reference's weights=None (bit-identical
        # results) and then does not stage it: a third less to upload per step, which is what eight ranks sharing one
        # host feel; the device-resident leg above keeps the per-pixel array the workload names
        bbe = BlendBatch(data["images"], psfs, data["model_psf"], data["centers"], data["bg_rms"],
                         weights=data["weights"], streams=args.e2e_streams, detect_unit_weights=True)
    # staging buffers allocated page-locked (cudaHostAlloc through torch) AFTER the rank was bound near its GPU, instead
    # of page-locking whatever pages the allocator had handed out before
    keep_pinned = []

    def pinned(a):
        if not isinstance(a, np.ndarray) or a.nbytes < 4096:
            return a
        t = torch.empty(a.nbytes, dtype=torch.uint8, pin_memory=True)            # any numpy dtype: a byte buffer, re-viewed
        v = t.numpy().view(a.dtype).reshape(a.shape)
        v[...] = a
        keep_pinned.append(t)
        return v
    res = {k: pinned(v) for k, v in bbe._dev.alloc_results(MAX_ITER, morph=True).items()}
    for k in list(bbe.host.keys()):
        if k == "psfs":
            continue
        bbe.host[k] = pinned(bbe.host[k])
    def e2e_step():
        _, o = bbe.fit_host(MAX_ITER, E_REL, check_every=args.check_every, out=res, morph=True)
        # per-blend result records gathered over NCCL (north_star: NCCL only for results)
        return shard.gather_records(shard.result_records(o, comp_off))
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_s = allmax(time.perf_counter() - t0)
    e2e = dict(value=blends_total * args.steps / e2e_s, unit=UNIT, h2d_bytes_per_step=int(allsum(float(bbe.h2d_bytes))),
               d2h_bytes_per_step=int(allsum(float(bbe.d2h_bytes))), ms_per_step=e2e_s * 1e3 / args.steps,
               streams_per_gpu=args.e2e_streams,
               weights="array of ones recognised at construction (weights=None of observation.py:148-151): not uploaded"
               if getattr(bbe, "uniform_weights", False) else "per-pixel array uploaded every step")
    if bbe is not bb:
        bbe.close()

    # ---- strong scaling (BASELINE config 3 as worded: ONE batch of 10k blends sharded over the GPUs): the same
    # --strong-blends blends whatever N is; sorted by component count and dealt round-robin so every rank gets the
    # same mix; device-timed like `value`
    strong = None
    if world > 1 and args.strong_blends > 0:
        Ks = np.array([len(synth.draw_scene(i, C, NY, NX)[0]["re"]) for i in range(args.strong_blends)])
        mine = shard.deal_sorted(Ks, rank, world)
        ds = synth.make_blends(mine, C, NY, NX, psf_size=PSF)
        bs = BlendBatch(ds["images"], ds["psfs"], ds["model_psf"], ds["centers"], ds["bg_rms"], weights=ds["weights"])
        bs.snapshot()
        for _ in range(min(args.warmup, 2)):
            bs.restore()
            bs.fit_device(MAX_ITER, E_REL, check_every=args.check_every)
        barrier()
        ev0.record()
        for _ in range(args.steps):
            bs.restore()
            bs.fit_device(MAX_ITER, E_REL, check_every=args.check_every)
        ev1.record()
        barrier()
        ms_s = allmax(ev0.elapsed_time(ev1))
        strong = dict(value=args.strong_blends * args.steps / (ms_s / 1e3), unit=UNIT, total_blends=args.strong_blends,
                      blends_per_gpu=int(allsum(float(bs.B)) / world), ms_per_step=ms_s / args.steps,
                      sharding="sorted by component count, dealt round-robin (rank r takes sorted[r::N])")
        bs.close()

    # ---- roofline of each kernel: fixed-iteration pass (e_rel = 0: every blend active in every launch)
    # (chunked batch: one chunk after another, so a "launch" below is the chunk launches of one iteration together)
    bb.restore()
    bb.fit_device(8, 0.0, check_every=1000, serial=True)
    bb.restore()
    fx = bb.fit_device(args.roofline_iters, 0.0, check_every=1000, profile=True, serial=True)
    per_launch, it_bytes = alg_bytes(comp_off, weight_planes=0 if bb.uniform_weights else 1)
    names = ["scl_k_forward_adjoint", "scl_k_lipschitz", "scl_k_prox", "scl_k_converge"]
    kernels = []
    for j in range(3):
        t = fx["kernel_ms"][j] / fx["steps"]
        gbs = per_launch[j] / (t * 1e-3) / 1e9
        kernels.append(dict(name=names[j], ms_per_launch=t, alg_bytes_per_launch=per_launch[j], gbs=gbs,
                            frac=gbs / hbm_peak, share=fx["kernel_ms"][j] / sum(fx["kernel_ms"])))
    top = max(range(3), key=lambda j: fx["kernel_ms"][j])
    # ncu evidence per kernel (DRAM traffic per launch, issue-slot and fp32-pipe utilisation: SURVEY 8(d) "give both
    # numbers"), captured under ncu by tools/ncu_stamp.py and stamped with a hash of the kernel's sources: evidence
    # taken from other sources than the ones this library was built from is dropped, not printed
    evidence = kernel_evidence()
    for kk in kernels:
        ev = evidence.get(kk["name"])
        kk["issue_frac"] = ev["issue_frac"] if ev else None
        kk["fma_pipe_frac"] = ev["fma_pipe_frac"] if ev else None
        kk["dram_traffic_per_launch"] = ev["traffic"] if ev else None
        kk["ncu"] = ev["stamp"] if ev else "no ncu capture of the current sources (profiles/kernel_evidence.json)"
    traffic = kernels[top]["dram_traffic_per_launch"]
    roofline = dict(bound="hbm", kernel=names[top], achieved=kernels[top]["gbs"], peak=hbm_peak, unit="GB/s",
                    frac=kernels[top]["frac"], traffic=traffic, traffic_source=kernels[top]["ncu"], peak_source=peak_src,
                    how="CUDA events around each launch, %d fixed iterations with every blend active (e_rel=0); "
                        "algorithmic bytes per launch per DESIGN.md" % fx["steps"],
                    iteration=dict(ms=fx["device_ms"] / fx["steps"], alg_bytes=it_bytes,
                                   gbs=it_bytes / (fx["device_ms"] / fx["steps"] * 1e-3) / 1e9,
                                   frac=it_bytes / (fx["device_ms"] / fx["steps"] * 1e-3) / 1e9 / hbm_peak),
                    timed_region_share={names[j]: kms[j] / max(kms.sum(), 1e-9) for j in range(4)},
                    kernels=kernels)
    if rank == 0:
        line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=args.warmup,
                    ms_per_step=ms / args.steps, higher_is_better=True, scaling="weak", vs_baseline=None,
                    dtype="f32", data="synthetic",
                    config=dict(workload=WORKLOAD, blends_per_gpu=int(bb.B), components_per_gpu=int(bb.K),
                                max_iter=MAX_ITER, e_rel=E_REL, mean_iters_per_blend=iters_total / blends_total,
                                l2="inputs per GPU (%.1f GB) exceed the 126 MB L2" % (bb.h2d_bytes / 1e9),
                                streams_per_gpu=args.streams,
                                weights=("an array of ones: recognised at construction as the reference's weights=None "
                                         "(observation.py:148-151), not stored or uploaded; algorithmic bytes with C_w = 0")
                                if bb.uniform_weights else "per-pixel array",
                                psfs="one set per blend, matched on the device" if args.per_blend_psfs else "one shared set",
                                sharding="blend-parallel, rank r owns blends [r*B,(r+1)*B), no data-path collective",
                                input_gen_s=t_gen, source_init_s=t_init, failed_sources=bb.n_failed,
                                host_binding=binding),
                    pgm_iters_per_sec=iters_total * args.steps / (ms / 1e3),
                    e2e=e2e, gpu_launches=int(launches), roofline=roofline, clocks=clocks)
        if cpu is not None:
            line["cpu_baseline"] = cpu
        if strong is not None:
            line["config"]["strong_scaling"] = strong
        if world == 1 and args.config4_steps > 0:
            # the other single-GPU configuration with a bench line of its own (bench.py --workload config4), short
            try:
                bb.close()
                c4 = scene_line(args, args.config4_steps, 1, False)
                line["config4"] = dict(workload=c4["config"]["workload"], value=c4["value"], unit=c4["unit"],
                                       steps=c4["steps"], ms_per_step=c4["ms_per_step"],
                                       iters_per_fit=c4["config"]["iters_per_fit"], components=c4["config"]["components"],
                                       e2e=c4["e2e"], gpu_launches=c4["gpu_launches"],
                                       ms_per_iteration=c4["roofline"]["iteration"]["ms"],
                                       kernel_ms=c4["roofline"]["kernel_ms"], roofline_frac=c4["roofline"]["frac"])
            except Exception as exc:        # the headline line must not depend on the secondary workload
                line["config4"] = dict(error="%s: %s" % (type(exc).__name__, exc))
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_reference(args):
    """The reference's algorithm on the host cores: the CPU oracle (the reference itself cannot be imported or built
    here: autograd, proxmin and Eigen are absent -- DESIGN.md)."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    # every step is a bounded sample of the workload (blends per worker process); --steps / --warmup are honoured as
    # given and the sample shrinks instead, so that the whole run stays within a few minutes
    per_worker = args.ref_per_worker
    if per_worker <= 0:
        per_worker = int(min(24, max(3, round(72 / max(args.warmup + args.steps, 1)))))
    vals = []
    cpu = None
    for i in range(args.warmup + args.steps):
        r = cpu_baseline(per_worker=per_worker, first_index=i * 1000)
        if i >= args.warmup:
            vals.append(r)
        cpu = r
    # blends / (mean busy time of a core): what the cores deliver when the queue never runs dry -- the figure a
    # per-step wall clock approaches on a long run, and the one that favours the CPU on a short sample
    v = float(np.mean([r["balanced_value"] for r in vals]))
    cpu = dict(cpu, value=v, sample="%d steps (+%d warm-up) of %d blends each (%d per worker process, fresh blends every "
                                    "step); value = blends x cores / summed fit time; the slowest-worker wall-clock "
                                    "rate of the same samples is %.1f blends/s"
                                    % (args.steps, args.warmup, cpu["cores"] * per_worker, per_worker,
                                       float(np.mean([r["value"] for r in vals]))))
    n = int(os.environ.get("WORLD_SIZE", 1))
    print(json.dumps(dict(metric=METRIC, value=v, unit=UNIT, impl="reference", n_gpus=n, steps=args.steps,
                          warmup=args.warmup, ms_per_step=1e3 * float(np.mean([r["fit_wall_s"] for r in vals])),
                          higher_is_better=True, scaling="weak", vs_baseline=None,
                          dtype="f32", data="synthetic", config=dict(workload=WORKLOAD, max_iter=MAX_ITER, e_rel=E_REL),
                          cpu_baseline=cpu,
                          e2e=dict(value=v, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)))


# ------------------------------------------------------------------------------------------ config 4 (one big scene)
SCENE_WORKLOAD = ("config4: one synthetic LSST-like 6-band 512x512 scene, 200 sources (70 % two-component), 41x41 PSFs, "
                  "monotonic + symmetric constraints, Blend.fit(200, e_rel=1e-2)")


def scene_oracle_step_seconds(batch, extra, n_steps=1):
    """One (or a few) iterations of the CPU oracle on the same scene, single thread: seconds per iteration."""
    from oracle import ref_blend as rb
    C, Ny, Nx = batch["C"], batch["Ny"], batch["Nx"]
    frame = rb.Frame((C, Ny, Nx), psfs=extra["mpsf"], dtype=np.float32)
    obs = rb.Observation(extra["images"], psfs=extra["psfs"], weights=batch["weights"][0]).match(frame)
    cw = rb._centroid_weight(frame)
    srcs, k, K = [], 0, len(batch["sed"])
    while k < K:
        if extra["multi"][k]:
            m = object.__new__(rb.MultiComponentSource)
            m.frame, m.symmetric, m.monotonic, m.parent = frame, True, True, None
            m.pixel_center, m.shift, m.centroid_weight = tuple(batch["center"][k]), (0., 0.), cw
            m.components = [rb.Component(frame, batch["sed"][j], batch["morph"][j]) for j in (k, k + 1)]
            srcs.append(m)
            k += 2
        else:
            s_ = object.__new__(rb.ExtendedSource)
            rb.Component.__init__(s_, frame, batch["sed"][k], batch["morph"][k])
            s_.symmetric, s_.monotonic, s_.centroid_weight = True, True, cw
            s_.pixel_center, s_.shift = tuple(batch["center"][k]), (0., 0.)
            srcs.append(s_)
            k += 1
    blend = rb.Blend(srcs, obs)
    t0 = time.perf_counter()
    for _ in range(n_steps):
        blend.step(E_REL)
    return (time.perf_counter() - t0) / n_steps, blend.mse


def run_scene(args):
    """BASELINE config 4 on one GPU ("replicas only" for --gpus > 1): scenes/s of Blend.fit(200, 1e-2)."""
    import torch
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0))))
    line = scene_line(args, args.steps, args.warmup, not args.no_cpu, world, rank, dist)
    if line is not None:
        print(json.dumps(line))


def scene_line(args, steps, warmup, with_cpu, world=1, rank=0, dist=None):
    """The config-4 measurement (run_scene's line; also the `config4` entry of the default line at N = 1)."""
    import torch
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    import __graft_entry__
    if rank == 0:
        __graft_entry__.build()
    if world > 1:
        dist.barrier()
    from scarlet_b200 import synth, _native as nat
    t0 = time.perf_counter()
    batch, extra = synth.scene_arrays(args.scene_size, args.scene_sources, 6)
    t_gen = time.perf_counter() - t0
    K, N, Cb = len(batch["sed"]), batch["Ny"] * batch["Nx"], batch["C"]
    bt = nat.Batch(**batch)
    bt.snapshot()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    iters, launches = 0, 0
    for i in range(warmup + steps):
        if i == warmup:
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            ev0.record()
            iters, launches = 0, 0
        bt.restore()
        info = bt.fit(MAX_ITER, E_REL, check_every=args.check_every)
        iters += info["steps"]
        launches += info["launches"]
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    out = bt.download()
    # end to end: host arrays -> upload -> fit -> download of the fitted scene
    up = {k: v for k, v in batch.items() if k not in ("C", "Ny", "Nx", "ker_half")}
    t0 = time.perf_counter()
    for _ in range(steps):
        bt.upload(**up)
        bt.fit(MAX_ITER, E_REL, check_every=args.check_every)
        res = bt.download()
    e2e_s = (time.perf_counter() - t0) / steps
    d2h = sum(v.nbytes for v in res.values() if isinstance(v, np.ndarray))
    # per-kernel time of a fixed-iteration pass
    bt.restore()
    fx = bt.fit(args.roofline_iters, 0.0, profile=True)
    km = [v / fx["steps"] for v in fx["kernel_ms"]]
    alg_k3 = 4.0 * N * (2 + Cb) * K
    alg_it = 4.0 * N * (2 * K + 2 * Cb) + 4 * Cb * PSF * PSF + 8 * K * Cb
    roofline = dict(bound="hbm", kernel="scene constraint chain (scl_k_wide_step, _source_center, _prep, _sym_rows, "
                                        "_sym_cols, _finish)",
                    achieved=alg_k3 / (km[2] * 1e-3) / 1e9, peak=hbm_peak,
                    unit="GB/s", frac=alg_k3 / (km[2] * 1e-3) / 1e9 / hbm_peak, traffic=None,
                    how="CUDA events around each kernel group, %d fixed iterations (e_rel=0); the chain ends when the "
                        "longest monotonic sweep does (~1000 dependent levels of a frame-sized component on one SM), "
                        "see DESIGN.md scene mode" % fx["steps"],
                    iteration=dict(ms=fx["device_ms"] / fx["steps"], alg_bytes=alg_it,
                                   gbs=alg_it / (fx["device_ms"] / fx["steps"] * 1e-3) / 1e9),
                    kernel_ms=dict(render_conv_adjoint=km[0], lipschitz=km[1], constraints=km[2], converge=km[3]))
    bt.close()
    if rank != 0:
        return None
    line = dict(metric=METRIC, value=world * steps / (ms / 1e3), unit=UNIT, n_gpus=world, steps=steps,
                warmup=warmup, ms_per_step=ms / steps, higher_is_better=True, scaling="weak", vs_baseline=None,
                dtype="f32", data="synthetic",
                config=dict(workload=SCENE_WORKLOAD, size=args.scene_size, sources=args.scene_sources, components=K,
                            max_iter=MAX_ITER, e_rel=E_REL, iters_per_fit=iters / steps,
                            l2="state per iteration (%.0f MB of morphologies) exceeds the 126 MB L2" % (4e-6 * N * K),
                            sharding="replicas only: one scene is one unit", input_gen_s=t_gen),
                pgm_iters_per_sec=world * iters / (ms / 1e3),
                e2e=dict(value=world / e2e_s, unit=UNIT, h2d_bytes_per_step=int(bt.h2d_bytes), d2h_bytes_per_step=int(d2h),
                         ms_per_step=e2e_s * 1e3),
                gpu_launches=int(launches), roofline=roofline)
    if with_cpu:
        t_it, _ = scene_oracle_step_seconds(batch, extra)
        per_fit = iters / steps
        line["cpu_baseline"] = dict(value=1.0 / (t_it * per_fit), unit=UNIT, cores=1, kind="port",
                                    sample="1 iteration of the same scene with the CPU oracle (%.1f s), scaled to the "
                                           "%.0f iterations the GPU fit took" % (t_it, per_fit),
                                    iters_per_sec=1.0 / t_it)
    return line


# ------------------------------------------------------------------------------------------ config 5 (multi-resolution)
MULTIRES_WORKLOAD = ("config5: joint fits of a 3-band 128x128 observation (41x41 PSFs) and a 5-band 34x34 LowResObservation "
                     "(pixel scale 4, 15x15 PSFs, sub-pixel offset, unrotated), %d CombinedExtendedSources per blend, "
                     "Blend.fit(200, e_rel=1e-2), a batch of blends of one geometry through scarlet.fit_many")


def run_multires(args):
    """BASELINE config 5 on one GPU, built and fitted through the drop-in Python API (scarlet.Frame / Observation /
    LowResObservation / CombinedExtendedSource / Blend).  One step = Blend.fit(200, 1e-2) from the initial sources."""
    import copy
    import torch
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    if int(os.environ.get("RANK", 0)) != 0:
        return
    import __graft_entry__
    __graft_entry__.build()
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    from wcs_stub import make_pair            # affine WCS stand-in (astropy is not in this image)
    import scarlet_b200 as scarlet
    from scarlet_b200 import psf as spsf
    rng = np.random.default_rng(20260411 + 5)
    HY = HX = 128
    LY = LX = 34
    ns = args.multires_sources
    chan = ["h0", "h1", "h2", "l0", "l1", "l2", "l3", "l4"]
    whr, wlr = make_pair((HY, HX), (LY, LX), scale=4.0, lr_offset=(.3, -.2))
    mpsf = spsf.generate_psf_image(spsf.gaussian, (41, 41), sigma=.9).image.astype(np.float32)
    hr_psfs = np.array([spsf.generate_psf_image(spsf.gaussian, (41, 41), sigma=1.2 + .15 * b).image for b in range(3)],
                       dtype=np.float32)
    lr_psfs = np.array([spsf.generate_psf_image(spsf.gaussian, (15, 15), sigma=.8 + .05 * b).image for b in range(5)],
                       dtype=np.float32)
    t0 = time.perf_counter()
    frame = scarlet.Frame((8, HY, HX), wcs=whr, psfs=mpsf[None], channels=chan)
    centers = []
    while len(centers) < ns:
        c = rng.uniform(24, HY - 24, 2)
        if all(np.hypot(*(c - o)) >= 14 for o in centers):
            centers.append(c)
    yy, xx = np.indices((HY, HX))
    truth = np.zeros((8, HY, HX), np.float32)
    for c in centers:
        re, q, pa = rng.uniform(2, 4), rng.uniform(.6, 1), rng.uniform(0, np.pi)
        u = (xx - c[1]) * np.cos(pa) + (yy - c[0]) * np.sin(pa)
        v = (-(xx - c[1]) * np.sin(pa) + (yy - c[0]) * np.cos(pa)) / q
        sed = rng.dirichlet(np.ones(8)) * 10 ** rng.uniform(3, 4)
        truth += (sed[:, None, None] * np.exp(-np.hypot(u, v) / re)[None]).astype(np.float32)
    # images: this package's own forward models of the true scene + noise
    ohr = scarlet.Observation(np.zeros((3, HY, HX), np.float32), psfs=hr_psfs, wcs=whr, channels=chan[:3]).match(frame)
    olr = scarlet.LowResObservation(np.zeros((5, LY, LX), np.float32), wcs=wlr, psfs=lr_psfs, channels=chan[3:]).match(frame)
    hr_img = (ohr.render(truth) + rng.normal(0, 1, (3, HY, HX))).astype(np.float32)
    lr_img = rng.normal(0, 1, (5, LY, LX)).astype(np.float32)
    ys, xs = olr._window()
    lr_img[:, ys, xs] += olr._render(truth)
    ohr = scarlet.Observation(hr_img, psfs=hr_psfs, wcs=whr, channels=chan[:3]).match(frame)
    t1 = time.perf_counter()
    olr = scarlet.LowResObservation(lr_img, wcs=wlr, psfs=lr_psfs, channels=chan[3:]).match(frame)
    t_match = time.perf_counter() - t1
    bg = [np.ones(3, np.float32), np.ones(5, np.float32)]
    # B blends of one geometry: every blend has its own scene and images; the observations of all blends share the
    # matched difference kernels and the ONE resample-and-convolve operator of the geometry (335 MB), so the batch runs
    # its low-resolution planes as matrix products over the blends (csrc/k_lowres.cu: scl_k_lowres_gemm)
    B = max(1, args.multires_blends)

    def scene(i):
        r = np.random.default_rng(20260411 + 500 + i)
        cs = []
        while len(cs) < ns:
            c = r.uniform(24, HY - 24, 2)
            if all(np.hypot(*(c - o)) >= 14 for o in cs):
                cs.append(c)
        tr = np.zeros((8, HY, HX), np.float32)
        for c in cs:
            re, q, pa = r.uniform(2, 4), r.uniform(.6, 1), r.uniform(0, np.pi)
            u = (xx - c[1]) * np.cos(pa) + (yy - c[0]) * np.sin(pa)
            v = (-(xx - c[1]) * np.sin(pa) + (yy - c[0]) * np.cos(pa)) / q
            sed = r.dirichlet(np.ones(8)) * 10 ** r.uniform(3, 4)
            tr += (sed[:, None, None] * np.exp(-np.hypot(u, v) / re)[None]).astype(np.float32)
        o_hr, o_lr = copy.copy(ohr), copy.copy(olr)
        o_hr.images = (ohr.render(tr) + r.normal(0, 1, (3, HY, HX))).astype(np.float32)
        li = r.normal(0, 1, (5, LY, LX)).astype(np.float32)
        li[:, ys, xs] += olr._render(tr)
        o_lr.images = li
        return cs, o_hr, o_lr

    blends0 = []
    for i in range(B):
        cs, o_hr, o_lr = (centers, ohr, olr) if i == 0 else scene(i)
        srcs = [scarlet.CombinedExtendedSource(frame, (round(c[0]) - HY / 2, round(c[1]) - HX / 2), [o_hr, o_lr], bg, obs_idx=0)
                for c in cs]
        blends0.append((srcs, [o_hr, o_lr]))
    t_setup = time.perf_counter() - t0
    from scarlet_b200.blend import pack_blends
    ms_list, its, launches = [], [], 0
    e2e_list = []
    for i in range(args.warmup + args.steps):
        blends = []
        for srcs0, obs in blends0:
            srcs = [copy.copy(s) for s in srcs0]
            for s, s0 in zip(srcs, srcs0):
                s._sed, s._morph = s0._sed.copy(), s0._morph.copy()
            blends.append(scarlet.Blend(srcs, obs))
        t0 = time.perf_counter()
        infos = scarlet.fit_many(blends, MAX_ITER, E_REL)
        dt = time.perf_counter() - t0
        if i >= args.warmup:
            ms_list.append(sum(f["device_ms"] for f in infos)); its.append(float(np.mean([b.it for b in blends])))
            launches += sum(f["launches"] for f in infos)
            e2e_list.append(dt)
    blend = blends[0]
    packed = pack_blends(blends)
    h2d = sum(v.nbytes for v in packed.values() if isinstance(v, np.ndarray)) + sum(
        v.nbytes for v in (packed["lr"] or {}).values() if isinstance(v, np.ndarray))
    ms = float(np.mean(ms_list))
    K = sum(len(s) for s, _ in blends0)
    line = dict(
        metric=METRIC, value=B * 1e3 / ms, unit=UNIT, n_gpus=1, steps=args.steps, warmup=args.warmup, ms_per_step=ms,
        higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
        config=dict(workload=MULTIRES_WORKLOAD % ns, blends=B, components=K, max_iter=MAX_ITER, e_rel=E_REL,
                    iters_per_fit=float(np.mean(its)), lr_operator_bytes=int(olr._lr_op.nbytes),
                    lowres_match_s=t_match, setup_s=t_setup,
                    note="scene mode (128x128 frame + 41x41 PSF exceeds one convolution tile); the blends share one "
                         "geometry: the low-resolution operator is applied as a matrix product over the blends"),
        pgm_iters_per_sec=B * float(np.mean(its)) / (ms / 1e3),
        e2e=dict(value=B / float(np.mean(e2e_list)), unit=UNIT, h2d_bytes_per_step=int(h2d),
                 d2h_bytes_per_step=int(4 * K * (HY * HX + 8)), ms_per_step=1e3 * float(np.mean(e2e_list)),
                 note="through the Python surface: pack (host), upload incl. the operator, fit, download, unpack"),
        gpu_launches=int(launches), logL_first_last=[float(blend.mse[0]), float(blend.mse[-1])])
    if not args.no_cpu:
        line["cpu_baseline"] = multires_cpu_baseline()
    print(json.dumps(line))


def multires_cpu_baseline(n_gpu=64):
    """The CPU oracle on a multi-resolution fit, timed at the size at which its adjoint is tractable (the fixture of
    tests/golden/make_golden_lowres.py: 3-band 40x40 + 5-band 12x12, 3 CombinedExtendedSources; the oracle needs the
    dense matrix of `_render`, 1600 renders at that size, 16384 at 128x128), next to the device on the same scene, as
    a batch of `n_gpu` copies: a like-for-like ratio at that size, stated as such."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers
    from scarlet_b200 import _native as nat
    gold = np.load(os.path.join(ROOT, "tests", "golden", "ref_golden_lowres.npz"))
    frame, ohr, olr = helpers.lowres_scene(gold, "A_")
    srcs, blend = helpers.lowres_blend(gold, frame, ohr, olr)
    olr.dense_operator(frame.shape)                       # one-off, not timed (the reference gets it from autograd)
    packs = helpers.pack([blend] * n_gpu)
    t0 = time.perf_counter()
    blend.fit(MAX_ITER, E_REL)
    t_cpu = time.perf_counter() - t0
    bt = nat.Batch(**packs)
    bt.snapshot()
    bt.fit(MAX_ITER, E_REL)                               # warm-up
    bt.restore()
    info = bt.fit(MAX_ITER, E_REL)
    out = bt.download()
    bt.close()
    return dict(value=1.0 / t_cpu, unit=UNIT, cores=1, kind="port",
                sample="ONE fit of the fixture-size scene (3-band 40x40 + 5-band 12x12, 3 sources, %d iterations) with the "
                       "CPU oracle on one core: %.2f s; NOT the 128x128 workload of this line" % (blend.it, t_cpu),
                same_size_device=dict(value=n_gpu / (info["device_ms"] / 1e3), unit=UNIT, blends=n_gpu,
                                      iterations=int(out["n_iter"][0]),
                                      note="the same scene as a batch of %d copies on the device" % n_gpu))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--multires-sources", type=int, default=6)
    ap.add_argument("--multires-blends", type=int, default=128, help="config5: blends of one geometry in the batch")
    ap.add_argument("--workload", default="config3", choices=["config3", "config4", "config5"])
    ap.add_argument("--scene-size", type=int, default=512)
    ap.add_argument("--scene-sources", type=int, default=200)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--blends", type=int, default=10000, help="blends per GPU")
    ap.add_argument("--check-every", type=int, default=8)
    ap.add_argument("--streams", type=int, default=1, help="independently driven chunks per GPU (BlendBatch streams=)")
    ap.add_argument("--e2e-streams", type=int, default=4,
                    help="chunks per GPU of the end-to-end leg (upload / loop / download pipelined chunk by chunk)")
    ap.add_argument("--roofline-iters", type=int, default=40)
    ap.add_argument("--strong-blends", type=int, default=10000,
                    help="N > 1: also time ONE batch of this many blends sharded over the N GPUs (0 = skip)")
    ap.add_argument("--ref-per-worker", type=int, default=0,
                    help="--impl reference: blends per worker process and step (0 = sized from --steps / --warmup)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--config4-steps", type=int, default=3,
                    help="N = 1: fits of the config-4 scene reported as `config4` in the default line (0 = skip)")
    ap.add_argument("--per-blend-psfs", action="store_true",
                    help="config 3 with one observed PSF set per blend (B x C kernel spectra matched on the device and "
                         "kept in HBM) instead of one shared set")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "config4":
        run_scene(args)
    elif args.workload == "config5":
        run_multires(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
