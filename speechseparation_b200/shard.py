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
