"""Multi-GPU plumbing: the path shards by independent event stream (BASELINE.json
configs[3]), one process per GPU, no collective on the data path.
torch.distributed is used only for the barrier and for the whole-job
aggregate (sum of units, max of time over ranks) that rank 0 reports."""
from __future__ import annotations

import os


def rank_world():
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def stream_seed(base_seed: int, rank: int) -> int:
    """Rank 0 tracks the named stream; every further rank an independent stream
    of the same shape (seeds 2000 + rank)."""
    return base_seed if rank == 0 else 2000 + rank


def streams_of_rank(n_streams: int, rank: int, world: int) -> list:
    """Stream s -> GPU s mod G."""
    return [s for s in range(n_streams) if s % world == rank]


def aggregate(units: float, seconds: float, device=None):
    """(sum of units over ranks, max of seconds over ranks); identity when not
    distributed.  Works with nccl (device tensors) and gloo (CPU tensors)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(units), float(seconds)
    u = torch.tensor([float(units)], dtype=torch.float64, device=device)
    s = torch.tensor([float(seconds)], dtype=torch.float64, device=device)
    dist.all_reduce(u, op=dist.ReduceOp.SUM)
    dist.all_reduce(s, op=dist.ReduceOp.MAX)
    return u.item(), s.item()


def gather_rows(row, device=None):
    """Every rank's row of floats, in rank order (a list of lists); [row] when not
    distributed.  Only used for the report: per-rank times and counts."""
    import torch
    import torch.distributed as dist
    row = [float(v) for v in row]
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [row]
    mine = torch.tensor(row, dtype=torch.float64, device=device)
    out = [torch.empty_like(mine) for _ in range(dist.get_world_size())]
    dist.all_gather(out, mine)
    return [t.tolist() for t in out]
