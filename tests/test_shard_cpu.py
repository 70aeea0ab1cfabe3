"""N > 1 host logic on CPU: sharding and the per-blend record gather over gloo, world_size 2."""
import os
import socket

import numpy as np
import torch.multiprocessing as mp

from scarlet_b200 import shard


def test_shard_range_partitions():
    for n, w in [(10, 3), (7, 8), (10000, 8), (0, 2)]:
        got = [i for r in range(w) for i in shard.shard_range(n, r, w)]
        assert got == list(range(n))
        sizes = [len(shard.shard_range(n, r, w)) for r in range(w)]
        assert max(sizes) - min(sizes) <= 1
    assert list(shard.weak_range(5, 2)) == [10, 11, 12, 13, 14]


def test_result_records():
    out = dict(n_iter=np.array([3, 0, 2]), active=np.array([0, 1, 0]),
               logL=np.array([[5., 4., 3.5, 0.], [0., 0., 0., 0.], [9., 8., 0., 0.]]))
    rec = shard.result_records(out, [0, 2, 5, 6])
    np.testing.assert_array_equal(rec, [[3, 1, 3.5, 2], [0, 0, 0., 3], [2, 1, 8., 1]])


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    own = list(shard.shard_range(7, rank, world))
    rec = np.array([[i, 1, 0.5 * i, 2] for i in own], dtype=np.float64)
    allrec = shard.gather_records(rec)
    q.put((rank, [a.tolist() for a in allrec]))
    dist.destroy_process_group()


def test_gather_records_gloo_world2():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    expect = [[[i, 1, 0.5 * i, 2] for i in shard.shard_range(7, r, 2)] for r in range(2)]
    assert res[0] == expect and res[1] == expect          # ragged shards (4 + 3 blends), same view on both ranks


def test_deal_sorted_partitions_and_balances():
    """Strong scaling: units sorted by size and dealt round-robin -- a partition, with per-rank totals within one
    largest unit of each other."""
    from scarlet_b200 import shard
    rng = np.random.default_rng(0)
    sizes = rng.integers(2, 9, 1000)
    parts = [shard.deal_sorted(sizes, r, 8) for r in range(8)]
    assert sorted(np.concatenate(parts).tolist()) == list(range(1000))
    tot = [int(sizes[p].sum()) for p in parts]
    assert max(tot) - min(tot) <= 8
    assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
