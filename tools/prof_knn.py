"""Profiling driver: one remove_duplicate -> index -> kNN(k)+features pass on a synthetic city cloud."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pointcloudprocess_b200 as pcp
from pointcloudprocess_b200 import sharded

n = int(sys.argv[1]) if len(sys.argv) > 1 else 5_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
ctx = pcp.Context(0)
job = sharded.ShardedPipeline(ctx, 0, 1, n, seed=5, leaf=0.02, k=20, torch=torch)
job.generate()
ctx.profile(True)
for _ in range(reps):
    m = job.step_device()
ctx.synchronize()
for s in ctx.profile_get():
    print("%-28s %9.3f ms  %d" % (s["name"], s["ms"] / reps, s["launches"]))
print("featured", m)
