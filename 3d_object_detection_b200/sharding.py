"""Frame-wise sharding across GPUs: frames are independent, so rank r of G owns a subset of
the frames of every batch and runs its own PillarFrontEnd replica; there is no collective on the data path.
Only timings / counters are reduced, over whatever process group the launcher set up (NCCL on GPUs, gloo in
the CPU tests)."""
from __future__ import annotations


def shard_frames(n_frames: int, rank: int, world_size: int, mode: str = "block"):
    """Indices of the frames rank `rank` owns.  "block": contiguous ceil-split; "cyclic": b -> b mod G."""
    if world_size < 1 or not 0 <= rank < world_size:
        raise ValueError(f"bad rank/world_size {rank}/{world_size}")
    if mode == "cyclic":
        return list(range(rank, n_frames, world_size))
    if mode != "block":
        raise ValueError(f"unknown sharding mode {mode!r}")
    base, extra = divmod(n_frames, world_size)
    lo = rank * base + min(rank, extra)
    return list(range(lo, lo + base + (1 if rank < extra else 0)))


def reduce_max_time(seconds: float, device=None) -> float:
    """MAX over ranks of a per-rank duration (the multi-GPU timing rule); identity without a process group."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return float(seconds)
    t = torch.tensor([seconds], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def reduce_sum(values, device=None):
    """SUM over ranks of a list of counters (frames, points, bytes)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return [float(v) for v in values]
    t = torch.tensor(list(values), dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return [float(v) for v in t.tolist()]
