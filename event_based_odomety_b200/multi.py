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


class FeatureShard:
    """One stream's features over the ranks of a job (north_star "partitioned by feature batch"):
    every rank feeds its own tracker context the SAME images and event batches; rank r tracks the patches
    whose track id = r (mod world) (ebt_set_shard) and, after every push, the ranks gather their patch
    states on the HOST (a gloo group: no NCCL collective, no GPU-to-GPU traffic) and merge them
    (ebt_merge_patch_state), so that the next image's association, window sizes and archive decisions are
    the same on every rank and the same as on one GPU.  Wraps a cabi.Tracker; everything else passes through."""

    def __init__(self, tracker, rank, world, group=None):
        self.tr, self.rank, self.world, self.group = tracker, rank, world, group
        tracker.set_shard(rank, world)

    def __getattr__(self, k):
        return getattr(self.tr, k)

    def exchange(self):
        import numpy as np
        import torch
        import torch.distributed as dist
        mine = self.tr.patch_state()
        if self.world == 1 or len(mine) == 0:
            return
        t = torch.from_numpy(np.ascontiguousarray(mine).reshape(-1))
        out = [torch.empty_like(t) for _ in range(self.world)]
        dist.all_gather(out, t, group=self.group)  # CPU tensors on a gloo group: a host-side gather
        self.tr.merge_patch_state([o.numpy().reshape(mine.shape) for o in out])

    def push_events(self, events):
        n = self.tr.push_events(events)
        self.exchange()
        return n

    def push_events_ptr(self, ptr, n_events):
        n = self.tr.push_events_ptr(ptr, n_events)
        self.exchange()
        return n

    def push_events_device(self, ptr, n_events):
        n = self.tr.push_events_device(ptr, n_events)
        self.exchange()
        return n


def host_group():
    """A gloo process group next to the default one (NCCL on the GPU box) for host-side gathers; the default
    group itself when that already is gloo; None when not distributed."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return None
    if dist.get_backend() == "gloo":
        return dist.group.WORLD
    return dist.new_group(backend="gloo")
