import sys, os, time
sys.path.insert(0, "/root/repo")
import pointcloudprocess_b200 as pcp
ctx = pcp.Context(0)
n = int(os.environ.get("PCP_RADIUS_N", 20_000_000))
cloud, props = ctx.alloc(n * 48), ctx.alloc(n * 48)
ctx.synth_cloud("forest", n, seed=3, out=cloud.ptr)
ix = pcp.Index(ctx, (cloud.ptr, n, 48))
for rep in range(3):
    ctx.profile(True); ctx.profile_reset()
    t = time.perf_counter()
    ix.features_radius(0.5, props=props.ptr, want_extra=False)
    ctx.synchronize()
    dt = (time.perf_counter() - t) * 1e3
    print(rep, "%.1f ms" % dt, {s["name"]: round(s["ms"], 2) for s in ctx.profile_get() if s["name"].startswith("radius")})
print("h", ix.cell_size())
