#!/usr/bin/env python
"""bench.py -- kNN(k=20)+feature points/sec on B200 (BASELINE.json metric), one process per GPU.

A step is one pass of the hot path over one cloud shard: PointCloudHelper::remove_duplicate (voxel grid,
leaf 0.02) -> KdTreeFLANN::setInputCloud (search structure) -> calculate_plan_parameter (kNN k=20 +
plane fit per point), i.e. BASELINE.json configs[4] ("full pipeline voxel_grid 0.02 m then kNN k=20
features on a 200M-point synthetic city-block cloud"), of which every GPU holds a 25 M-point spatial
shard (200 M / 8): weak scaling.

  value  : points that received features per second, inputs resident in HBM, device timed (CUDA events)
  e2e    : same through the C ABI with HOST buffers (pinned), H2D of the cloud and D2H of the
           LAS_POINT_PROPERTY records inside the timed region
  roofline: dominant kernel (knn_fast_kernel): 72 B/point algorithmic bytes / its CUDA-event duration
  cpu_baseline: the oracle (CPU restatement of the reference path) on a bounded sample, host cores

`--impl reference` times the oracle on the host cores only (rank 0), never touching the CUDA library.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "kNN(k=20)+feature points/sec"
UNIT = "points/s"
LEAF = 0.02
K = 20
SHARD_POINTS = 25_000_000   # 200 M / 8
ALGO_BYTES_PER_POINT = 72.0  # xyz doubles in (24) + LAS_POINT_PROPERTY out (48); SURVEY 8(d)


# ---- numpy restatement of the `city` generator (csrc/synth.cu) for the CPU arm ----------------------------
def _splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15))
    x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


class _Rng:
    def __init__(self, seed, stream, i):
        with np.errstate(over="ignore"):
            self.base = (np.uint64(seed) * np.uint64(0x9E3779B97F4A7C15) + np.uint64(stream) * np.uint64(0xD1B54A32D192ED03)
                         + i.astype(np.uint64) * np.uint64(16))

    def u(self, c):
        with np.errstate(over="ignore"):
            return (_splitmix64(self.base + np.uint64(c)) >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)

    def gauss(self, c):
        a = np.maximum(self.u(c), 1e-300)
        return np.sqrt(-2.0 * np.log(a)) * np.cos(6.283185307179586 * self.u(c + 1))


def synth_city_numpy(n_total, seed, centred=False):
    """Same formulas as synth_kernel kind 3 (city); libm rounding may differ from CUDA's in the last bit."""
    i = np.arange(n_total, dtype=np.uint64)
    r = _Rng(seed, 0, i)
    nt = float(n_total)
    side = max(600.0 * np.sqrt(nt / 2e8), 20.0)
    m = r.u(0)
    x = np.zeros(n_total)
    y = np.zeros(n_total)
    z = np.zeros(n_total)
    g = m < 0.40
    x[g], y[g], z[g] = (side * r.u(1))[g], (side * r.u(2))[g], (0.02 * r.gauss(3))[g]
    b = (m >= 0.40) & (m < 0.85)
    nbox = int(max(120.0 * nt / 2e8, 1.0))
    bi = np.minimum((r.u(1) * nbox).astype(np.uint64), np.uint64(nbox - 1))
    br = _Rng(seed, 9, bi)
    fx = np.minimum(15.0 + 45.0 * br.u(0), side * 0.5)
    fy = np.minimum(15.0 + 45.0 * br.u(1), side * 0.5)
    hh = 6.0 + 54.0 * br.u(2)
    bx, by = br.u(3) * (side - fx), br.u(4) * (side - fy)
    a1, a2, a3 = fx * hh, fy * hh, fx * fy
    f = r.u(2) * (2 * a1 + 2 * a2 + a3)
    s, t, off = r.u(3), r.u(4), 0.01 * r.gauss(5)
    w1 = b & (f < a1)
    w2 = b & (f >= a1) & (f < 2 * a1)
    w3 = b & (f >= 2 * a1) & (f < 2 * a1 + a2)
    w4 = b & (f >= 2 * a1 + a2) & (f < 2 * a1 + 2 * a2)
    w5 = b & (f >= 2 * a1 + 2 * a2)
    for w, xx, yy, zz in ((w1, bx + s * fx, by + off, t * hh), (w2, bx + s * fx, by + fy + off, t * hh),
                          (w3, bx + off, by + s * fy, t * hh), (w4, bx + fx + off, by + s * fy, t * hh),
                          (w5, bx + s * fx, by + t * fy, hh + off)):
        x[w], y[w], z[w] = xx[w], yy[w], zz[w]
    tr = (m >= 0.85) & (m < 0.95)
    ntree = int(max(3000.0 * nt / 2e8, 1.0))
    ti = np.minimum((r.u(1) * ntree).astype(np.uint64), np.uint64(ntree - 1))
    trr = _Rng(seed, 7, ti)
    tx, ty = trr.u(0) * side, trr.u(1) * side
    rad, hgt, sig = 0.1 + 0.3 * trr.u(2), 8.0 + 17.0 * trr.u(3), 1.0 + trr.u(4)
    trunk = tr & (m < 0.88)
    crown = tr & (m >= 0.88)
    a = 6.283185307179586 * r.u(2)
    x[trunk], y[trunk], z[trunk] = (tx + rad * np.cos(a))[trunk], (ty + rad * np.sin(a))[trunk], (hgt * r.u(3))[trunk]
    x[crown], y[crown], z[crown] = (tx + sig * r.gauss(2))[crown], (ty + sig * r.gauss(4))[crown], \
        (hgt + sig * r.gauss(6))[crown]
    c = m >= 0.95
    x[c], y[c], z[c] = (side * r.u(1))[c], (side * r.u(2))[c], (60.0 * r.u(3))[c]
    if not centred:
        x += 483189.100569
        y += 3371741.645173
        z += 430.094726
    from oracle import oracle as orc
    cloud = np.zeros(n_total, dtype=orc.POINT_DTYPE)
    cloud["x"], cloud["y"], cloud["z"], cloud["w"] = x, y, z, 1.0
    with np.errstate(over="ignore"):
        cloud["rgba"] = (_splitmix64(i ^ (np.uint64(seed) << np.uint64(32))) & np.uint64(0xffffff)).astype(np.uint32)
    cloud["stamp_id"] = (i % np.uint64(10)).astype(np.uint32)
    return cloud


# ---- CPU arm ------------------------------------------------------------------------------------------------
def cpu_pipeline_once(orc, cloud, threads):
    """remove_duplicate (single thread, std::sort, as voxel_grid.h:948) -> kd-tree (leaf 15) -> OpenMP loop of
    nearestKSearch(20) + plane fit (calculate_feature.cpp:222-233, 119-206).  Returns (seconds, featured points)."""
    t0 = time.perf_counter()
    dd, _ = orc.remove_duplicate(cloud, LEAF)
    orc.features_knn(dd, K, want_eig=False, want_lps=False, threads=threads)
    return time.perf_counter() - t0, len(dd)


def cpu_sample_points(orc, threads, budget_s=15.0):
    probe = synth_city_numpy(100_000, 5)
    t, m = cpu_pipeline_once(orc, probe, threads)
    rate = m / t
    n = int(min(max(rate * budget_s, 200_000), 4_000_000))
    return n


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import oracle as orc
    orc.lib()
    threads = orc.lib().orc_num_threads()
    n = args.sample if args.sample else cpu_sample_points(orc, threads, 8.0)
    cloud = synth_city_numpy(n, 5)
    for _ in range(args.warmup if args.warmup < 2 else 1):
        cpu_pipeline_once(orc, cloud, threads)
    times, m = [], 0
    for _ in range(args.steps):
        t, m = cpu_pipeline_once(orc, cloud, threads)
        times.append(t)
    ms = 1e3 * sum(times) / len(times)
    value = m / (ms * 1e-3)
    sample = "city cloud n_total=%d (same generator/density as the GPU shard), leaf %.2f -> %d points, k=%d" % (n, LEAF, m, K)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "configs[4] pipeline on a bounded CPU sample: " + sample},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ---- clocks --------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop, self.th = index, [], threading.Event(), None

    def _run(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                p = [s.strip() for s in out.strip().split(",")]
                if len(p) >= 6:
                    self.rows.append(p)
            except Exception:
                pass
            self.stop.wait(0.2)

    def __enter__(self):
        self.th = threading.Thread(target=self._run, daemon=True)
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.th.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for j, n in enumerate(names) if any(r[2 + j].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows)}


# ---- GPU arm -------------------------------------------------------------------------------------------------
def run_gpu(args):
    import torch
    import torch.distributed as dist
    import pointcloudprocess_b200 as pcp
    from pointcloudprocess_b200 import sharded

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = pcp.Context(local)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)

    n_shard = args.points
    comm = None
    if world > 1:  # payloads over NCCL, message sizes host to host over gloo
        comm = sharded.DistComm(dist, torch, rank, world, torch.device("cuda", local),
                                host_group=dist.new_group(backend="gloo"))
    job = sharded.ShardedPipeline(ctx, rank, world, n_shard, seed=5, leaf=LEAF, k=K, comm=comm, torch=torch)
    job.generate()  # this rank's spatial shard, resident in HBM (untimed)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- device-resident timing ----
    for _ in range(args.warmup):
        job.step_device()
    ctx.profile(True)
    ctx.profile_reset()
    l0 = ctx.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    with ClockSampler(local) as clk:
        e0.record(stream)
        featured = 0
        for _ in range(args.steps):
            featured += job.step_device()
        e1.record(stream)
        barrier()
    ms_local = e0.elapsed_time(e1)
    ms_total = max_over_ranks(ms_local)
    launches = ctx.launch_count() - l0
    stages = ctx.profile_get()
    ctx.profile(False)
    per_rank = None
    if world > 1:  # every rank's own clock and stage timers, for the report (host to host)
        mine = {"ms_per_step": ms_local / args.steps, "stages": {st["name"]: round(st["ms"] / args.steps, 3) for st in stages},
                "halo": job.halo_info()}
        per_rank = [None] * world
        dist.all_gather_object(per_rank, mine, group=comm.host_group)
    total_featured = sum_over_ranks(float(featured))
    ms_per_step = ms_total / args.steps
    value = total_featured / (ms_total * 1e-3)

    # ---- roofline of the dominant kernel (rank 0's stage timers: CUDA events on the launching stream) ----
    roof = None
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md 6.65 TB/s)"
    kst = [s for s in stages if s["name"] == "knn.fast_kernel"]
    if kst and kst[0]["ms"] > 0:
        s = kst[0]
        achieved = s["bytes"] / (s["ms"] * 1e-3) / 1e9
        traffic = None
        issue = None
        try:
            cap = json.load(open(os.path.join(ROOT, "profiles", "knn_fast_traffic.json")))
            traffic = cap.get("dram_bytes_per_launch")
            issue = cap.get("issue_active_pct")
        except Exception:
            pass
        roof = {"bound": "hbm", "kernel": "knn_fast_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "kernel_ms_per_launch": s["ms"] / s["launches"], "algorithmic_bytes_per_launch": s["bytes"] / s["launches"],
                "kernel_share_of_step": s["ms"] / (ms_total if world == 1 else e0.elapsed_time(e1)),
                # what actually limits the kernel (ncu capture under profiles/): issue slots busy, % of peak
                "limiter": {"what": "instruction issue", "issue_slots_busy_pct": issue,
                            "source": "profiles/knn_fast_traffic.json (ncu --set full)"}}

    # ---- end to end through the C ABI with host buffers ----
    e2e_ms, h2d, d2h, e2e_feat = job.time_e2e(max(1, args.steps), barrier, torch.cuda.synchronize)
    e2e_ms = max_over_ranks(e2e_ms)
    e2e_total = sum_over_ranks(float(e2e_feat))
    e2e_value = e2e_total / (e2e_ms * 1e-3)

    line = None
    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu:
            from oracle import oracle as orc
            orc.lib()
            threads = orc.lib().orc_num_threads()
            n = cpu_sample_points(orc, threads, 15.0)
            cloud = synth_city_numpy(n, 5)
            t, m = cpu_pipeline_once(orc, cloud, threads)
            cpu = {"value": m / t, "unit": UNIT, "cores": threads, "kind": "port",
                   "sample": "city cloud n_total=%d (same generator/density), leaf %.2f -> %d points, k=%d, %.1f s"
                             % (n, LEAF, m, K, t)}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": "configs[4]: city cloud, %d points per GPU (%d total), remove_duplicate leaf %.2f "
                                       "-> index build -> kNN k=%d + plane-fit features; inputs (%.1f GB/GPU) larger "
                                       "than L2, no flush" % (n_shard, n_shard * world, LEAF, K, n_shard * 48 / 1e9),
                           "points_per_gpu": n_shard, "featured_points_per_step": total_featured / args.steps,
                           "parallelism": "spatial XY tiles + halo, 1 rank/GPU" if world > 1 else "single GPU",
                           "halo": job.halo_info()},
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "ms_per_step": e2e_ms,
                        "how": "pinned host cloud -> C ABI -> pinned host records, wall clock over all steps; upload of "
                               "step i+1 and download of step i-1 overlap step i on the library's copy lanes"},
                "gpu_launches": int(launches), "clocks": clk.summary(), "roofline": roof, "cpu_baseline": cpu,
                "fallback_queries_per_step": sum(s["bytes"] for s in stages if s["name"] == "knn.warp_kernel") / 72.0 / args.steps,
                "stages_ms_per_step": {s["name"]: round(s["ms"] / args.steps, 4) for s in stages}}
        if per_rank:
            names = sorted({k for r in per_rank for k in r["stages"]})
            line["per_rank_ms_per_step"] = [round(r["ms_per_step"], 3) for r in per_rank]
            line["stages_ms_per_step_max_over_ranks"] = {k: max(r["stages"].get(k, 0.0) for r in per_rank) for k in names}
            line["boundary_queries_per_rank"] = [(r["halo"] or {}).get("boundary_queries") for r in per_rank]
        print(json.dumps(line), flush=True)
    job.close()
    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--points", type=int, default=SHARD_POINTS, help="points per GPU")
    ap.add_argument("--sample", type=int, default=0, help="reference arm: CPU sample size (0 = sized to a few seconds)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
