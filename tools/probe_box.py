#!/usr/bin/env python
"""What the GPU box looks like and how far the single-device pipeline goes: host memory, NUMA nodes, PCIe copy rates
(pinned, both directions, alone and together), then the full pipeline (remove_duplicate 0.02 -> index -> kNN k=20 +
plane fit) on city clouds of growing size up to configs[4]'s 200 M points, with the stage timers.
    python tools/probe_box.py [sizes in millions ...]
"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def sh(cmd):
    try:
        return subprocess.run(cmd, shell=True, capture_output=True, text=True, timeout=60).stdout.strip()
    except Exception as e:  # noqa: BLE001
        return "failed: %s" % e


def main():
    sizes = [int(a) for a in sys.argv[1:]] or [25, 100, 200]
    print("== host ==")
    print(sh("nproc; free -g | head -2; ls /sys/devices/system/node | grep node; "
             "for n in /sys/devices/system/node/node*; do echo $n $(cat $n/cpulist) $(grep MemTotal $n/meminfo); done"))
    print(sh("nvidia-smi topo -m | head -20"))
    print(sh("nvidia-smi --query-gpu=index,name,memory.total,pci.bus_id --format=csv"))
    import torch
    import pointcloudprocess_b200 as pcp
    from pointcloudprocess_b200 import sharded
    torch.cuda.set_device(0)
    ctx = pcp.Context(0)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)

    print("== copies (pinned, 1.2 GB) ==")
    nb = 1_200_000_000
    h1, p1 = ctx.pinned(nb)
    h2, p2 = ctx.pinned(nb)
    h1[:] = 1
    h2[:] = 2
    d1, d2 = ctx.alloc(nb), ctx.alloc(nb)
    for name, ops in (("h2d", [(0, d1.ptr, p1)]), ("d2h", [(1, p2, d2.ptr)]), ("both", [(0, d1.ptr, p1), (1, p2, d2.ptr)])):
        for rep in range(3):
            ctx.synchronize()
            t0 = time.perf_counter()
            for lane, dst, src in ops:
                ctx.copy_async(lane, dst, src, nb)
            for lane, _, _ in ops:
                ctx.copy_wait(lane)
            dt = time.perf_counter() - t0
        print("%s: %.1f GB/s per direction (%.1f ms)" % (name, nb / dt / 1e9, dt * 1e3))
    ctx.free_pinned(p1)
    ctx.free_pinned(p2)
    d1.free()
    d2.free()

    for mpts in sizes:
        n = mpts * 1_000_000
        print("== pipeline, %d M points ==" % mpts, flush=True)
        try:
            job = sharded.ShardedPipeline(ctx, 0, 1, n, seed=5, leaf=0.02, k=20, torch=torch)
            job.generate()
            for it in range(3):
                if it == 2:
                    ctx.profile(True)
                    ctx.profile_reset()
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                m = job.step_device()
                torch.cuda.synchronize()
                dt = time.perf_counter() - t0
                print("  step %d: %d -> %d points, %.1f ms, %.1f M points/s, mem %.1f GB" %
                      (it, n, m, dt * 1e3, m / dt / 1e6, torch.cuda.mem_get_info()[0] / 1e9), flush=True)
            st = ctx.profile_get()
            ctx.profile(False)
            print("  stages:", json.dumps({s["name"]: round(s["ms"], 2) for s in st}))
            if os.environ.get("PCP_PROBE_DEBUG"):
                os.environ["PCP_KNN_DEBUG"] = "1"
                job.step_device()
                del os.environ["PCP_KNN_DEBUG"]
            job.close()
            del job
            ctx.trim()
            torch.cuda.empty_cache()
        except Exception as e:  # noqa: BLE001
            print("  FAILED: %r" % (e,), flush=True)
            break
    ctx.close()


if __name__ == "__main__":
    main()
