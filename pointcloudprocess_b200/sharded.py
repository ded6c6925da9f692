"""One shard of the voxel-grid -> kNN -> feature pipeline per GPU: the host-side driver that
bench.py and the multi-GPU tests run, one process per GPU.

The cloud is cut into slabs of voxel columns along x, the way the reference chunks its work into 5 m
trajectory blocks (point_cloud_spliter.cpp:221-234) and CloudGrid keys 1 m tiles (cloud_grid.cpp:34-78):
every voxel belongs to exactly one rank, so remove_duplicate needs no exchange; the kNN stage needs the
voxelised points within `halo` metres of the slab faces from the two neighbouring ranks (one NCCL
send/recv pair per face), and queries whose k-th neighbour is not provably inside slab+halo are answered
exactly with candidates from the neighbour (second, small exchange).

All compute is libpcp_cuda.so; torch is used for device buffers, streams and torch.distributed only.
"""
import ctypes as C

import numpy as np

from . import api


class ShardedPipeline:
    def __init__(self, ctx, rank, world, n_per_rank, seed=5, leaf=0.02, k=20, kind="city", comm=None, torch=None,
                 halo=1.0):
        self.ctx, self.rank, self.world = ctx, rank, world
        self.n_per_rank, self.seed, self.leaf, self.k, self.kind = int(n_per_rank), seed, leaf, k, kind
        self.comm, self.torch, self.halo = comm, torch, float(halo)
        self.cloud = None       # DeviceArray: this rank's input points (48 B records), resident in HBM
        self.n = 0
        self.work = None        # DeviceArray: working copy that remove_duplicate filters in place
        self.props = None       # DeviceArray: LAS_POINT_PROPERTY records
        self._pinned = []
        self._halo_stats = {}

    # ---- setup (untimed) ----
    def generate(self):
        if self.world != 1:
            raise NotImplementedError("multi-rank shards: see ShardedPipeline.generate_slab")
        ctx = self.ctx
        self.n = self.n_per_rank
        self.cloud = ctx.alloc(self.n * 48)
        ctx.synth_cloud(self.kind, self.n, seed=self.seed, out=self.cloud)
        self.work = ctx.alloc(self.n * 48)
        self.props = ctx.alloc(self.n * 48)
        ctx.synchronize()

    def halo_info(self):
        return self._halo_stats or None

    # ---- one step, inputs resident in HBM ----
    def step_device(self):
        """remove_duplicate(leaf) -> index build -> kNN(k)+features.  Returns the number of featured points."""
        ctx = self.ctx
        ctx.memcpy_async(self.work, self.cloud, self.n * 48)  # remove_duplicate is in place: keep the input
        m = ctx.remove_duplicate((self.work, self.n, 48), self.leaf)
        ix = api.Index(ctx, (self.work, m, 48))
        try:
            ix.features_knn(self.k, props=self.props, want_eig=False, want_lps=False)
        finally:
            ix.free()
        return m

    # ---- end to end: host buffers through the C ABI ----
    def time_e2e(self, steps, barrier):
        """Times `steps` passes whose input cloud starts in pinned HOST memory and whose LAS_POINT_PROPERTY
        records end in pinned HOST memory.  Returns (ms_per_step, h2d_bytes, d2h_bytes, featured_total/steps)."""
        import time
        ctx = self.ctx
        host_in, pin_in = ctx.pinned(self.n * 48)
        host_props, pin_props = ctx.pinned(self.n * 48)
        ctx.memcpy(pin_in, self.cloud, self.n * 48)
        ctx.synchronize()
        t_ms, feat, d2h = [], 0, 0
        torch = self.torch
        for it in range(steps + 1):
            barrier()
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            stream = torch.cuda.current_stream()
            t0 = time.perf_counter()
            e0.record(stream)
            m = self._step_host(pin_in, pin_props)
            e1.record(stream)
            barrier()
            wall = (time.perf_counter() - t0) * 1e3
            if it == 0:
                continue  # first pass warms the staging arena
            # host-side work (ctypes calls, synchronisations) sits between the events: take the larger clock
            t_ms.append(max(e0.elapsed_time(e1), wall))
            feat += m
            d2h = m * 48
        ctx.free_pinned(pin_in)
        ctx.free_pinned(pin_props)
        return sum(t_ms) / len(t_ms), self.n * 48, d2h, feat / len(t_ms)

    def _step_host(self, pin_in, pin_props):
        ctx = self.ctx
        # H2D of the cloud (host pointer handed to the C ABI), filtered on the device, features to host
        ctx.memcpy_async(self.work, pin_in, self.n * 48)
        m = ctx.remove_duplicate((self.work, self.n, 48), self.leaf)
        ix = api.Index(ctx, (self.work, m, 48))
        try:
            ix.features_knn(self.k, props=self.props, want_eig=False, want_lps=False)
        finally:
            ix.free()
        ctx.memcpy(pin_props, self.props, m * 48)
        return m

    def close(self):
        for a in (self.cloud, self.work, self.props):
            if a is not None:
                a.free()
        self.cloud = self.work = self.props = None
