#!/usr/bin/env python
"""Where the end-to-end pass goes: the device-resident step, the copy-only passes and the pipelined passes of
ShardedPipeline.time_e2e side by side on one GPU.   python tools/e2e_trace.py [million points] [steps]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import pointcloudprocess_b200 as pcp
    from pointcloudprocess_b200 import sharded
    n = (int(sys.argv[1]) if len(sys.argv) > 1 else 25) * 1_000_000
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    ctx = pcp.Context(0)
    job = sharded.ShardedPipeline(ctx, 0, 1, n, torch=torch)
    job.generate()
    for _ in range(3):
        job.step_device()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        job.step_device()
    torch.cuda.synchronize()
    print("device-resident step: %.2f ms" % ((time.perf_counter() - t0) * 1e3 / steps))
    nosync = lambda: None  # noqa: E731
    for label, kw in (("copy only", dict(copy_only=True)), ("pipelined e2e", {})):
        ms, h2d, d2h, units = job.time_e2e(steps, nosync, torch.cuda.synchronize, **kw)
        print("%s: %.2f ms per pass (%d passes; up %.2f GB, down %.2f GB)" % (label, ms, steps, h2d / 1e9, d2h / 1e9))
    ctx.close()


if __name__ == "__main__":
    main()
