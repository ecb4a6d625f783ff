"""Multi-GPU sharding of the separation path.

Every op of the model is independent along dim 0 (rows = clips / sessions; reference
bsrnn.py:61,391-395 with merge_channels=False) and streaming state is per row
(bsrnn.py:125-128), so the path shards by rows with NO data-path collective: one process per
GPU, weights replicated (30 MB), a contiguous block of rows each.  T (serial time LSTM) and
bands (coupled by the band LSTM) are never split.  The only communication is the start
barrier and the max-over-ranks of the device time, both outside the timed region.
"""
from __future__ import annotations

from typing import Tuple

import torch


def shard_range(n_rows: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous [lo, hi) block of rows for `rank`; sizes differ by at most one."""
    if world < 1 or not (0 <= rank < world) or n_rows < 0:
        raise ValueError(f"bad shard request rows={n_rows} rank={rank} world={world}")
    base, rem = divmod(n_rows, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def aggregate_throughput(local_units: float, local_seconds: float, device=None) -> Tuple[float, float]:
    """(sum of units over ranks, max of seconds over ranks) via torch.distributed when it is
    initialised (NCCL on GPUs, gloo in the CPU tests); identity for a single process."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(local_units), float(local_seconds)
    dev = device if device is not None else ("cuda" if dist.get_backend() == "nccl" else "cpu")
    u = torch.tensor([local_units], dtype=torch.float64, device=dev)
    t = torch.tensor([local_seconds], dtype=torch.float64, device=dev)
    dist.all_reduce(u, op=dist.ReduceOp.SUM)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(u.item()), float(t.item())


def bind_to_gpu_numa_node(device_index: int) -> int:
    """Pin this process to the CPUs of the NUMA node its GPU hangs off, BEFORE any pinned host buffer is allocated
    (pages are placed on the node of the allocating thread). With one process per GPU and host buffers of 113 MB per
    direction and step, copies that cross the socket interconnect are what keeps the host-buffer path from scaling.
    Returns the node, or -1 when the topology is not exposed (then nothing is changed)."""
    import os
    try:
        p = torch.cuda.get_device_properties(device_index)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read().strip())
        if node < 0:
            return -1
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return -1
        os.sched_setaffinity(0, cpus)
        return node
    except (OSError, ValueError, AttributeError):
        return -1
