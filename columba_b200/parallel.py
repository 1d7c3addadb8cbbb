"""Multi-GPU driver of the path: reads are sharded over the ranks (one process
per GPU, index replicated in every GPU's HBM), there is no data-path
collective; only the final counters and per-read hit counts are reduced
(reference: the worker threads of src/parallel.cpp:88-115 share one index and
add their Counters at the end, src/parallel.cpp:1198-1229).

The collectives go through torch.distributed, so the same code runs over NCCL
(GPU tensors, NVLink) and over gloo (CPU tensors, used by the tests).
"""
import numpy as np

from .api import COUNTER_NAMES, STAT_NAMES

BATCH_READS = 1 << 16  # >= 64 K reads per batch and GPU


def shard_batches(n_reads, rank, world, batch=BATCH_READS):
    """Batches [begin, end) of this rank: batches are dealt round-robin, so a
    repeat-rich stretch of the input does not end up on one GPU."""
    out = []
    for i, b in enumerate(range(0, n_reads, batch)):
        if i % world == rank:
            out.append((b, min(n_reads, b + batch)))
    return out


def bind_to_gpu_numa_node(device):
    """Restricts this process to the CPUs NVML reports as local to the GPU, so
    that the page-locked read / result buffers allocated afterwards (first
    touch) live on the GPU's NUMA node and its copies do not cross the socket
    interconnect.  With several GPUs per host this is what keeps the end-to-end
    rate from being bound by the host.  Returns the CPU list, or None when the
    affinity is unknown or outside the CPUs this process may use."""
    import os
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        try:
            h = pynvml.nvmlDeviceGetHandleByUUID("GPU-" + str(torch.cuda.get_device_properties(device).uuid))
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(device)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return sorted(cpus)
    except Exception:
        return None


def _dist():
    import torch.distributed as dist
    return dist if dist.is_available() and dist.is_initialized() else None


def _device_for_collectives():
    import torch
    dist = _dist()
    if dist is not None and dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def reduce_counters(counters, stats=None):
    """Sum of the 17 reference counters (and the device statistics) over all
    ranks; every rank gets the totals."""
    import torch
    vals = [int(counters[k]) for k in COUNTER_NAMES] + [int((stats or {}).get(k, 0)) for k in STAT_NAMES]
    t = torch.tensor(vals, dtype=torch.int64, device=_device_for_collectives())
    dist = _dist()
    if dist is not None and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    v = t.cpu().tolist()
    n = len(COUNTER_NAMES)
    return dict(zip(COUNTER_NAMES, v[:n])), dict(zip(STAT_NAMES, v[n:]))


def gather_hit_counts(n_reads, local):
    """local: {global read id -> hit count} as two arrays (ids, counts).
    Returns the per-read hit counts of the whole input on every rank (4 B per
    read; ranks own disjoint reads, so a sum is a gather)."""
    import torch
    ids, counts = local
    t = torch.zeros(n_reads, dtype=torch.int32, device=_device_for_collectives())
    if len(ids):
        t[torch.as_tensor(np.asarray(ids, np.int64), device=t.device)] = torch.as_tensor(
            np.asarray(counts, np.int32), device=t.device)
    dist = _dist()
    if dist is not None and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.cpu().numpy()


class ShardedMapper:
    """Maps this rank's share of a read set with one Mapper (one GPU)."""

    def __init__(self, mapper, rank=0, world=1, batch=BATCH_READS):
        self.mapper, self.rank, self.world, self.batch = mapper, rank, world, batch

    def map_reads(self, seqs, max_dist, min_dist=0):
        """seqs: list of read strings (the whole input, identical on all ranks).
        Returns (results, counters, stats, hit_counts): results = list of
        (global read offset, BatchResult) of this rank; counters / stats / hit
        counts are global (reduced over the ranks)."""
        from .api import Mapper
        results = []
        ctr = dict.fromkeys(COUNTER_NAMES, 0)
        st = dict.fromkeys(STAT_NAMES, 0)
        ids, counts = [], []
        for b, e in shard_batches(len(seqs), self.rank, self.world, self.batch):
            reads, off = Mapper.pack(seqs[b:e])
            r = self.mapper.map_batch(reads, off, max_dist, min_dist)
            results.append((b, r))
            for k in COUNTER_NAMES:
                ctr[k] += r.counters[k]
            for k in STAT_NAMES:
                st[k] += r.stats[k]
            ids.append(np.arange(b, e))
            counts.append(r.read_count)
        ids = np.concatenate(ids) if ids else np.zeros(0, np.int64)
        counts = np.concatenate(counts) if counts else np.zeros(0, np.int32)
        gctr, gst = reduce_counters(ctr, st)
        hits = gather_hit_counts(len(seqs), (ids, counts))
        return results, gctr, gst, hits
