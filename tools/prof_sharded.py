"""All ranks of the sharded pipeline in lock step on ONE GPU (debugging aid): per-stage times summed over ranks."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pointcloudprocess_b200 as pcp
from pointcloudprocess_b200 import sharded

world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n_total = int(sys.argv[2]) if len(sys.argv) > 2 else 25_000_000 * world
ctx = pcp.Context(0)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
jobs = [sharded.ShardedPipeline(ctx, r, world, n_total, seed=5, leaf=0.02, k=20, torch=torch) for r in range(world)]
for j in jobs:
    j.generate()
ctx.profile(True)
t0 = time.perf_counter()
ms = sharded.run_lockstep([j._step_gen(j._p(j.src)) for j in jobs])
torch.cuda.synchronize()
print("plan bounds", jobs[0].plan.bounds, "range sizes", [j.n for j in jobs])
print("lock-step pass over %d ranks: %.1f ms" % (world, (time.perf_counter() - t0) * 1e3))
for s in ctx.profile_get():
    if s["ms"] > 1.0:
        print("%-28s %9.3f ms  %d" % (s["name"], s["ms"], s["launches"]))
print([j.stats for j in jobs])
# warm passes: wall clock against the sum of the library's stages (the rest is the driver: torch ops, syncs)
for rep in range(3):
    ctx.profile_reset()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    sharded.run_lockstep([j._step_gen(j._p(j.src)) for j in jobs])
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) * 1e3
    st = ctx.profile_get()
    print("warm pass %d: %.1f ms wall over %d ranks, library stages %.1f ms" % (rep, wall, world, sum(s["ms"] for s in st)))
