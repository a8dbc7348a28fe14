"""Sharding of a batch by series id across the GPUs of one box.

Series are independent and the roll-up is per service, so rank r scores a
contiguous range of series cut at service boundaries on its own GPU / ctx, and
the only "collective" is a host gather of the verdict records.  No data-path
NCCL: `torch.distributed` (gloo or nccl) is used solely to gather the small
result arrays when the ranks are separate processes.
"""
import numpy as np


def shard_bounds(n_series, world, service_of_series=None):
    """[lo, hi) per rank: balanced contiguous ranges; with `service_of_series`
    (non-decreasing) the cuts are moved forward to the next service boundary so
    that no service is split across ranks."""
    cuts = [0]
    for r in range(1, world):
        c = (n_series * r) // world
        if service_of_series is not None:
            s = np.asarray(service_of_series)
            while 0 < c < n_series and s[c] == s[c - 1]:
                c += 1
        cuts.append(max(c, cuts[-1]))
    cuts.append(n_series)
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def score_sharded(score_fn, rank, world, n_series, service_of_series=None, gather=None):
    """Run `score_fn(lo, hi) -> structured ndarray of (hi - lo) records` on this
    rank's range and gather all ranges on every rank.

    gather: callable(list_slot, local_bytes ndarray) -> list of per-rank uint8
    arrays (e.g. built on torch.distributed.all_gather_object); None = single
    process (world must be 1)."""
    lo, hi = shard_bounds(n_series, world, service_of_series)[rank]
    local = score_fn(lo, hi)
    if world == 1:
        return local
    parts = gather(np.frombuffer(local.tobytes(), np.uint8))
    return np.concatenate([np.frombuffer(p.tobytes(), local.dtype) for p in parts])


def torch_gather():
    """gather() over the default torch.distributed process group (gloo or nccl)."""
    import torch.distributed as dist

    def g(local_u8):
        out = [None] * dist.get_world_size()
        dist.all_gather_object(out, local_u8)
        return out
    return g


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def bind_host_to_gpu(pci_bus_id, sysfs="/sys/bus/pci/devices"):
    """Pin the calling process to the CPUs of the NUMA node its GPU hangs off, so that the pinned
    staging buffers it allocates afterwards (first touch) and its copy threads sit next to that
    GPU's PCIe root.  With 8 ranks each pushing 6 GB per batch, rows that cross the socket
    interconnect are what limits the end-to-end rate.

    pci_bus_id: "0000:1b:00.0" (any case).  Returns the CPU set, or None when the box does not
    say (single node, no sysfs): then nothing is changed."""
    import os
    dev = os.path.join(sysfs, pci_bus_id.lower())
    try:
        with open(os.path.join(dev, "numa_node")) as f:
            node = int(f.read().strip())
        if node < 0:
            return None
        with open(os.path.join(dev, "local_cpulist")) as f:
            cpus = _parse_cpulist(f.read())
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        if not cpus or cpus == allowed:
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except (OSError, ValueError, AttributeError):
        return None
