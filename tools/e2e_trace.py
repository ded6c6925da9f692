"""Per-pass wall-clock trace of the pipelined host-to-host path (debugging aid)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pointcloudprocess_b200 as pcp
from pointcloudprocess_b200 import sharded

n = int(sys.argv[1]) if len(sys.argv) > 1 else 25_000_000
ctx = pcp.Context(0)
job = sharded.ShardedPipeline(ctx, 0, 1, n, seed=5, leaf=0.02, k=20)
job.generate()
for _ in range(2):
    job.step_device()
ctx.synchronize()
t0 = time.perf_counter()
for _ in range(3):
    job.step_device()
ctx.synchronize()
print("device pass %.1f ms" % ((time.perf_counter() - t0) / 3 * 1e3))
nb = n * 48
hin, pin = ctx.pinned(nb)
hout, pout = ctx.pinned(nb)
dev = ctx.alloc(nb)
ctx.memcpy(pin, job._p(job.cloud), nb)
for name, fn in (("upload", lambda: ctx.copy_async(0, dev.ptr, pin, nb)), ("download", lambda: ctx.copy_async(1, pout, dev.ptr, nb))):
    for rep in range(2):
        ctx.synchronize(); t0 = time.perf_counter(); fn(); ctx.copy_wait(0); ctx.copy_wait(1)
        print("%s alone %.1f ms (%.1f GB/s)" % (name, (time.perf_counter() - t0) * 1e3, nb / (time.perf_counter() - t0) / 1e9))
ctx.synchronize(); t0 = time.perf_counter(); ctx.copy_async(0, dev.ptr, pin, nb); ctx.copy_async(1, pout, dev.ptr, nb); ctx.copy_wait(0); ctx.copy_wait(1)
print("both directions %.1f ms" % ((time.perf_counter() - t0) * 1e3))
ctx.synchronize(); t0 = time.perf_counter(); ctx.copy_async(0, dev.ptr, pin, nb); job.step_device(); ctx.synchronize(); ctx.copy_wait(0)
print("pass + concurrent upload %.1f ms" % ((time.perf_counter() - t0) * 1e3))
ctx.synchronize(); t0 = time.perf_counter(); ctx.copy_async(0, dev.ptr, pin, nb); ctx.copy_async(1, pout, dev.ptr, nb); job.step_device(); ctx.synchronize(); ctx.copy_wait(0); ctx.copy_wait(1)
print("pass + concurrent upload + download %.1f ms" % ((time.perf_counter() - t0) * 1e3))
for steps in (5, 10):
    ms = job.time_e2e(steps, lambda: None, ctx.synchronize)[0]
    print("pipelined %d passes: %.1f ms per pass" % (steps, ms))
