"""Device-resident timings of the other BASELINE.json configs (not bench lines): configs[1] voxel grid on the 10 M strip,
configs[2] radius features on the 20 M forest plot, configs[3] k=32 features on a 25 M city shard."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pointcloudprocess_b200 as pcp

which = sys.argv[1:] or ["voxel", "radius", "k32"]
ctx = pcp.Context(0)


def timed(fn, reps=3):
    fn(); ctx.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        r = fn()
    ctx.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3, r


def stages():
    return {s["name"]: round(s["ms"], 3) for s in ctx.profile_get()}


if "voxel" in which:
    n = 10_000_000
    cloud, work = ctx.alloc(n * 48), ctx.alloc(n * 48)
    ctx.synth_cloud("strip", n, seed=2, out=cloud.ptr)

    def f():
        ctx.memcpy_async(work.ptr, cloud.ptr, n * 48)
        return ctx.remove_duplicate((work.ptr, n, 48), 0.05)
    ms, m = timed(f)
    print("configs[1] remove_duplicate leaf 0.05, strip %d -> %d points: %.2f ms, %.1f M pts/s, algorithmic %.0f GB/s"
          % (n, m, ms, n / ms / 1e3, (32.0 * n + 32.0 * m) / ms / 1e6))
    cloud.free(); work.free()
if "radius" in which:
    n = int(os.environ.get("PCP_RADIUS_N", 20_000_000))
    cloud, props = ctx.alloc(n * 48), ctx.alloc(n * 48)
    ctx.synth_cloud("forest", n, seed=3, out=cloud.ptr)
    ctx.profile(True)

    def f():
        ix = pcp.Index(ctx, (cloud.ptr, n, 48))
        ix.features_radius(0.5, props=props.ptr, want_extra=False)
        ix.free()
        return n
    ms, _ = timed(f, 1)
    print("configs[2] radius 0.5 features, forest %d points: %.1f ms, %.2f M pts/s" % (n, ms, n / ms / 1e3))
    print("   ", {k: v for k, v in stages().items() if k.startswith("radius")})
    ctx.profile(False)
    cloud.free(); props.free()
if "k32" in which:
    n = 25_000_000
    cloud, props = ctx.alloc(n * 48), ctx.alloc(n * 48)
    ctx.synth_cloud("city", n, seed=4, out=cloud.ptr)
    ctx.profile(True)

    def f():
        ix = pcp.Index(ctx, (cloud.ptr, n, 48))
        ix.features_knn(32, props=props.ptr, want_eig=False, want_lps=False)
        ix.free()
        return n
    ms, _ = timed(f, 2)
    print("configs[3] k=32 features, city %d points (one shard): %.1f ms, %.1f M pts/s" % (n, ms, n / ms / 1e3))
    print("   ", {k: round(v / 3, 2) for k, v in stages().items() if k.startswith("knn")})
