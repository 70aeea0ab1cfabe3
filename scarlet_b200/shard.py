"""Blend-parallel sharding across ranks (one process per GPU).

Blends are independent units (each reference `Blend` owns its components and observations, blend.py:24-46), so the
data path needs no collective: rank r fits its own blends.  The only exchange is the gather of one fixed-size result
record per blend after a fit, over `torch.distributed` (NCCL on GPUs, gloo in the CPU tests).
"""
import numpy as np

RECORD_FIELDS = ("n_iter", "converged", "logL_last", "n_comp")


def shard_range(n_total, rank, world):
    """Contiguous, balanced slice of `range(n_total)` owned by `rank` (sizes differ by at most one)."""
    base, extra = divmod(int(n_total), int(world))
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


def deal_sorted(sizes, rank, world):
    """Strong-scaling layout: unit indices sorted by size (stable) and dealt round-robin, so every rank
    gets the same mix of small and large units; rank r owns sorted[r::world]."""
    order = np.argsort(np.asarray(sizes), kind="stable")
    return order[int(rank)::int(world)]


def weak_range(per_rank, rank):
    """Weak-scaling layout used by bench.py: rank r owns blends [r*B, (r+1)*B)."""
    return range(rank * int(per_rank), (rank + 1) * int(per_rank))


def result_records(out, comp_off):
    """float64 [B, 4] records (RECORD_FIELDS) from a `_native.Batch.download()` dict."""
    n = np.asarray(out["n_iter"], dtype=np.int64)
    B = len(n)
    last = np.zeros(B)
    if out.get("logL") is not None and out["logL"].shape[1] > 0:
        last = out["logL"][np.arange(B), np.clip(n - 1, 0, out["logL"].shape[1] - 1)]
        last = np.where(n > 0, last, 0.0)
    return np.stack([n.astype(np.float64), (np.asarray(out["active"]) == 0).astype(np.float64), last,
                     np.diff(np.asarray(comp_off)).astype(np.float64)], axis=1)


def gather_records(records, device=None):
    """All-gather per-blend records of every rank; returns the list of per-rank arrays (rank order).
    Ranks may own different numbers of blends.  Without an initialised process group: [records]."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [np.asarray(records)]
    world = dist.get_world_size()
    dev = device if device is not None else ("cuda" if dist.get_backend() == "nccl" else "cpu")
    rec = torch.as_tensor(np.ascontiguousarray(records, dtype=np.float64), device=dev)
    counts = torch.zeros(world, dtype=torch.int64, device=dev)
    counts[dist.get_rank()] = rec.shape[0]
    dist.all_reduce(counts)
    nmax = int(counts.max().item())
    pad = torch.zeros((nmax, rec.shape[1]), dtype=torch.float64, device=dev)
    pad[:rec.shape[0]] = rec
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad)
    return [b[:int(c)].cpu().numpy() for b, c in zip(bufs, counts.tolist())]
