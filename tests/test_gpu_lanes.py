"""Copy lanes (pcp_copy_async / pcp_copy_fence / pcp_copy_wait) and the pipelined host-to-host path built on them."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_copy_lanes_order_against_compute(ctx):
    import pointcloudprocess_b200 as pcp
    cloud = ctx.synth_cloud("city", 200000, seed=3)
    ref = ctx.remove_duplicate(cloud, 0.05)
    n, nb = len(cloud), len(cloud) * 48
    hin, pin = ctx.pinned(nb)
    hout, pout = ctx.pinned(nb)
    dev = ctx.alloc(nb)
    try:
        for rep in range(3):  # the same buffers again and again: every reuse is ordered by the lanes alone
            hin[:] = np.frombuffer(cloud.tobytes(), dtype=np.uint8)
            hout[:] = 0
            ctx.copy_async(0, dev.ptr, pin, nb)
            ctx.copy_fence(0)
            m = ctx.remove_duplicate((dev.ptr, n, 48), 0.05)
            ctx.copy_async(1, pout, dev.ptr, m * 48)
            ctx.copy_wait(1)
            got = np.frombuffer(hout[:m * 48].tobytes(), dtype=pcp.POINT_DTYPE)
            assert m == len(ref)
            for f in ("x", "y", "z", "w", "rgba", "stamp_id"):
                assert got[f].tobytes() == ref[f].tobytes(), f
        with pytest.raises(pcp.PcpError):
            ctx.copy_fence(7)
    finally:
        dev.free()
        ctx.free_pinned(pin)
        ctx.free_pinned(pout)


def test_pipelined_host_to_host_equals_device_pass(ctx):
    from pointcloudprocess_b200 import sharded
    job = sharded.ShardedPipeline(ctx, 0, 1, 300000, seed=9, leaf=0.02, k=20)
    job.generate()
    try:
        m = job.step_device()
        ref = np.empty(m * 48, dtype=np.uint8)
        ctx.memcpy(ref.ctypes.data, job._p(job.props), m * 48)
        keep = []
        ms, h2d, d2h, feat = job.time_e2e(5, lambda: None, ctx.synchronize, keep=keep)
        assert feat == m and h2d == 300000 * 48 and d2h == m * 48 and ms > 0
        assert len(keep) == 2
        for rec in keep:  # the last two passes went through different buffer sets
            assert rec.tobytes() == ref.tobytes()
    finally:
        job.close()
