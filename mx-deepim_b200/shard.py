"""Multi-GPU plumbing of the path: pairs are independent, so a batch is block-partitioned over
ranks with NO data-path collective (the reference likewise treats GPUs as independent slices,
/root/reference/lib/pair_matching/batch_updater_py_multi.py:176-365).  The only collectives are
the optional scalar reductions of loss / metric sums (PointMatchingLossMetric,
/root/reference/deepim/core/metric.py:108-120) and the max-over-ranks of a timing."""
from __future__ import annotations

import torch
import torch.distributed as dist


def partition(num_pairs: int, world_size: int, rank: int):
    """Contiguous block [lo, hi) of rank `rank`; sizes differ by at most one."""
    base, rem = divmod(int(num_pairs), int(world_size))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def rank_seed(base_seed: int, rank: int, set_index: int = 0) -> int:
    """Seed of a rank's synthetic input set (one seed per rank = base + rank)."""
    return int(base_seed) + 1000 * int(rank) + int(set_index)


def _is_dist():
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


def reduce_max(value: float, device=None) -> float:
    """Max over ranks of a host scalar (timings are reported as the slowest rank)."""
    if not _is_dist():
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_scalar(value: float, device=None):
    """The host scalar of every rank, in rank order (e.g. each rank's own device time, so that
    the spread behind a max-over-ranks figure is visible)."""
    if not _is_dist():
        return [float(value)]
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return [float(o.item()) for o in out]


def allreduce_metric(loss_sum: torch.Tensor, count: int):
    """Sum over ranks of (sum of the per-pair loss sums, number of samples): the
    PointMatchingLossMetric numerator / denominator.  8-16 bytes, latency bound."""
    t = torch.stack([loss_sum.double().sum(),
                     torch.tensor(float(count), dtype=torch.float64, device=loss_sum.device)])
    if _is_dist():
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t[0].item()), float(t[1].item())


def bind_to_gpu_numa(device_index: int):
    """Pin this process to the CPU cores NVML lists as local to GPU `device_index`, so that the
    pinned staging buffers it allocates afterwards (first touch) sit on the GPU's own NUMA node
    and the host feed of eight ranks does not cross the socket interconnect.  Best effort:
    returns the core list, or None when NVML / the affinity call is unavailable."""
    import os

    try:
        import pynvml

        pynvml.nvmlInit()
        try:
            h = None
            try:  # NVML enumerates in PCI order; CUDA's index may differ (CUDA_VISIBLE_DEVICES)
                pr = torch.cuda.get_device_properties(int(device_index))
                bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
                h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
            except Exception:
                h = None
            if h is None:
                h = pynvml.nvmlDeviceGetHandleByIndex(int(device_index))
            n_words = (os.cpu_count() + 63) // 64
            words = pynvml.nvmlDeviceGetCpuAffinity(h, n_words)
        finally:
            pynvml.nvmlShutdown()
        cores = [64 * w + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1]
        allowed = sorted(set(cores) & set(os.sched_getaffinity(0)))
        if not allowed:
            return None
        os.sched_setaffinity(0, allowed)
        return allowed
    except Exception:
        return None
