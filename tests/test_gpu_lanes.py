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
    """Odd and even pass counts end in different buffer sets: both deliver the device-resident pass's records."""
    import torch
    import pointcloudprocess_b200 as pcp
    from pointcloudprocess_b200 import sharded
    try:
        job = sharded.ShardedPipeline(ctx, 0, 1, 300000, seed=9, leaf=0.02, k=20, torch=torch)
        job.generate()
        m = job.step_device()
        ref = np.zeros(m, dtype=pcp.PACKED_PROP_DTYPE)
        ctx.props_pack(job._p(job.props), m, ref)
        ctx.synchronize()
        for steps in (4, 5):
            keep = []
            ms, h2d, d2h, feat = job.time_e2e(steps, lambda: None, torch.cuda.synchronize, keep=keep)
            assert feat == m and h2d == 300000 * 32 and d2h == m * 24 and ms > 0
            assert keep[0].tobytes() == ref.tobytes()
        job.close()
    finally:
        torch.cuda.synchronize()
        ctx.set_stream(None)
