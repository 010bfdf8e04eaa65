"""Tiny case for compute-sanitizer (racecheck / memcheck): 12 batched problems + a 2-frame 240x180 stream."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from event_based_odomety_b200 import synth
from event_based_odomety_b200.cabi import Tracker

ps = synth.make_problems(12, 240, 180, seed=11)
tr = Tracker(240, 180)
tr.add_frame(ps.frames[0])
sol, nab = tr.optimize_batch(ps.problems, ps.events)
tr.close()
S = synth.make_stream(240, 180, 0.10, 300e3, seed=5, max_corners=12)
g = Tracker(240, 180, 100)
g.set_products(1)
n = synth.run_stream(g, S, 3)
print("sanitize case: batch", len(sol), "stream optimisations", n)
