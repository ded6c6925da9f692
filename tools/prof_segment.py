"""Region growing at scale: stage times and loop statistics (not a bench line)."""
import sys, time
import numpy as np
sys.path.insert(0, "/root/repo")
import pointcloudprocess_b200 as pcp

n = int(sys.argv[1]) if len(sys.argv) > 1 else 5_000_000
kind = sys.argv[2] if len(sys.argv) > 2 else "city"
ctx = pcp.Context(0)
cloud = ctx.synth_cloud(kind, n, seed=4)
cloud = ctx.remove_duplicate(cloud, 0.04)   # point_cloud_spliter.cpp:127 runs remove_duplicate(0.04) first
print("points", len(cloud))
ix = pcp.Index(ctx, cloud)
for which in ("knn", "rpca"):
    props = ix.features_knn(20, want_eig=False, want_lps=False)["props"] if which == "knn" else ix.features_rpca(20)
    for rep in range(2):
        ctx.profile(True); ctx.profile_reset()
        t = time.time()
        r = ix.region_growing(props.copy(), want_stats=True)
        dt = time.time() - t
        st = {s["name"]: round(s["ms"], 2) for s in ctx.profile_get() if s["name"].startswith(("segment", "knn"))}
        ctx.profile(False)
    sizes = np.diff(r["offsets"])
    print(which, "wall %.3f s" % dt, "segments", len(sizes), "largest", sorted(sizes)[-3:], "members", len(r["points"]), r["stats"], st)
