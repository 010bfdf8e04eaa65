"""N>1 path on the CPU: two gloo ranks, each tracking its own independent
stream (sharding by stream, no data-path collective), aggregated as bench.py
does (sum of units, max of time)."""
import os
import socket
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from event_based_odomety_b200 import multi, synth
    from tests._oracle import CpuTracker, load_oracle
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    assert multi.rank_world() == (rank, world, rank)
    S = synth.make_stream(120, 90, 0.10, 80e3, seed=multi.stream_seed(1001, rank), max_corners=12)
    t = CpuTracker(load_oracle(), 120, 90, 20)
    n = synth.run_stream(t, S)
    units, secs = multi.aggregate(n, 1.0 + rank)
    rows = multi.gather_rows([rank, n, 1.0 + rank])  # bench.py's per-rank report
    assert [r[0] for r in rows] == list(range(world)) and rows[rank][1] == n
    assert sum(r[1] for r in rows) == units and max(r[2] for r in rows) == secs
    out.put((rank, n, units, secs, len(S.events)))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_shard_by_stream_and_aggregate():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    res = sorted(q.get(timeout=240) for _ in ps)
    [p.join(timeout=60) for p in ps]
    assert all(p.exitcode == 0 for p in ps)
    (r0, n0, u0, s0, e0), (r1, n1, u1, s1, e1) = res
    assert n0 > 0 and n1 > 0 and e0 != e1            # different streams per rank
    assert u0 == u1 == n0 + n1 and s0 == s1 == 2.0   # sum of units, max of time


def test_stream_placement():
    from event_based_odomety_b200 import multi
    for world in (1, 2, 4, 8):
        got = sorted(s for r in range(world) for s in multi.streams_of_rank(64, r, world))
        assert got == list(range(64))
        assert max(len(multi.streams_of_rank(64, r, world)) for r in range(world)) == 64 // world
    assert multi.stream_seed(1002, 0) == 1002 and multi.stream_seed(1002, 3) == 2003
    assert multi.aggregate(5, 2.0) == (5.0, 2.0)
    assert multi.gather_rows([1, 2.5]) == [[1.0, 2.5]]
