"""One stream's features sharded over two ranks (north_star: "partitioned across the GPUs by feature batch
... with only a host-side gather of tracks"): two processes, each with its own tracker context fed the same
images and event batches, rank r tracking the patches with track id = r (mod 2) and exchanging patch state
on the host after every push (event_based_odomety_b200/multi.FeatureShard, gloo group).  The union of the
two ranks' optimisation records must equal the records of the same stream tracked by one context: same
optimisations at the same stream events, bit-identical poses (the same kernel code ran on the same inputs).
Both ranks use cuda:0 here (the test box has one GPU); bench.py --shard features puts them on separate GPUs."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _stream():
    from event_based_odomety_b200 import synth
    return synth.make_stream(346, 260, 0.14, 2e6, seed=1072, max_corners=400)


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from event_based_odomety_b200 import multi, synth
    from event_based_odomety_b200.cabi import Tracker
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    S = _stream()
    t = multi.FeatureShard(Tracker(S.width, S.height, 400, device=0), rank, world, multi.host_group())
    n = synth.run_stream(t, S)
    rec = t.records()
    tracks = t.tracks()
    arch = t.tracks(True)
    out.put((rank, n, rec.tobytes(), tracks.tobytes(), arch.tobytes()))
    dist.barrier()
    t.close()
    dist.destroy_process_group()


@pytest.mark.gpu
def test_two_feature_shards_equal_one_context():
    from event_based_odomety_b200 import synth
    from event_based_odomety_b200.cabi import Tracker
    from event_based_odomety_b200.records import OPT_RECORD, TRACK
    S = _stream()
    one = Tracker(S.width, S.height, 400)
    n_one = synth.run_stream(one, S)
    want, want_tracks, want_arch = one.records().copy(), one.tracks().copy(), one.tracks(True).copy()
    one.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    res = sorted(q.get(timeout=600) for _ in ps)
    [p.join(timeout=120) for p in ps]
    assert all(p.exitcode == 0 for p in ps)
    recs = [np.frombuffer(r[2], OPT_RECORD) for r in res]
    assert res[0][1] + res[1][1] == n_one > 2000 and min(res[0][1], res[1][1]) > 0.3 * n_one
    for r, rec in enumerate(recs):
        assert (rec["track_id"] % 2 == r).all()          # every rank optimised only its own tracks
    got = np.concatenate(recs)
    got = got[np.lexsort((got["track_id"], got["event_index"]))]  # reference order: by event, then list position
    assert got.tobytes() == want.tobytes()
    # after the last exchange both ranks hold the full, identical track list — the one-context list
    for r in res:
        tr = np.frombuffer(r[3], TRACK)
        assert np.array_equal(tr["track_id"], want_tracks["track_id"])
        for k in ("lost", "init", "num_of_events", "n_optimizations", "rect_x", "rect_y", "flow_dir", "current_ts"):
            assert np.array_equal(tr[k], want_tracks[k]), k
        ar = np.frombuffer(r[4], TRACK)
        assert np.array_equal(ar["track_id"], want_arch["track_id"])
