"""N > 1 host logic on CPU: world_size-2 gloo run of the partition / seed / reduction helpers."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from mx_deepim_b200 import shard

    lo, hi = shard.partition(257, world, rank)
    mx = shard.reduce_max(10.0 + rank)
    s, n = shard.allreduce_metric(torch.tensor([1.0 + rank, 2.0]), hi - lo)
    every = shard.gather_scalar(2.5 * (rank + 1))
    assert every == [2.5, 5.0], every  # each rank's own value, in rank order
    q.put((rank, lo, hi, mx, s, n, shard.rank_seed(100, rank, 2)))
    dist.barrier()
    dist.destroy_process_group()


def test_partition_covers_batch():
    from mx_deepim_b200 import shard

    for B in (1, 16, 256, 1024, 1023):
        for W in (1, 2, 4, 8):
            spans = [shard.partition(B, W, r) for r in range(W)]
            assert spans[0][0] == 0 and spans[-1][1] == B
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    assert shard.reduce_max(3.5) == 3.5  # no process group: identity
    assert shard.gather_scalar(3.5) == [3.5]


def test_gloo_world_size_2():
    import sys
    from conftest import ROOT

    assert ROOT in sys.path
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29000 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, lo0, hi0, mx0, s0, n0, seed0), (r1, lo1, hi1, mx1, s1, n1, seed1) = res
    assert (lo0, hi0, lo1, hi1) == (0, 129, 129, 257)
    assert mx0 == mx1 == 11.0
    assert s0 == s1 == (1.0 + 2.0) + (2.0 + 2.0) and n0 == n1 == 257.0
    assert (seed0, seed1) == (102, 1102)


def test_bind_to_gpu_numa_is_best_effort():
    """No NVML / no GPU here: the helper reports None and leaves the affinity alone."""
    import os

    from mx_deepim_b200 import shard

    before = os.sched_getaffinity(0)
    assert shard.bind_to_gpu_numa(0) is None
    assert os.sched_getaffinity(0) == before
