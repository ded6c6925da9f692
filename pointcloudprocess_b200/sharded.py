"""One shard of the voxel-grid -> kNN -> feature pipeline per GPU: the host-side driver that
bench.py and the multi-GPU tests run, one process per GPU.

The cloud is cut into slabs of voxel columns along x, the way the reference chunks its work into 5 m
trajectory blocks (point_cloud_spliter.cpp:221-234) and CloudGrid keys 1 m tiles (cloud_grid.cpp:34-78):
  * every shard is centred on the serial-order centroid of the WHOLE cloud and voxelised in the whole cloud's
    grid box, and slabs are cut on voxel-column boundaries, so every voxel belongs to exactly one rank and
    remove_duplicate reproduces the single-device cells, centroids and order with no exchange;
  * the kNN stage needs the voxelised points within `halo` metres of the slab faces from the two neighbouring
    ranks (exchange 1: point records);
  * a query whose k-th neighbour distance exceeds its distance to the outer face of the halo is answered
    exactly in a second, small exchange: the neighbour returns its own k nearest candidates for that query and
    the owner merges the lists (pcp_features_merge).
All compute is libpcp_cuda.so; torch is used for device buffers, streams and torch.distributed only.

The pipeline is written as a generator that yields at every exchange point, so the same code runs under
torch.distributed (NCCL / gloo) and, for single-GPU parity tests, with all ranks emulated in one process.
"""
import ctypes as C
import math
import os
import sys

import numpy as np

from . import api

DBL_MIN = 2.2250738585072014e-308


class ShardPlan:
    """What every rank must agree on: centre, grid box, slab boundaries (voxel columns along x)."""

    def __init__(self, centre, count, min_b, div_b, leaf, bounds, shard_counts, halo):
        self.centre = np.asarray(centre, dtype=np.float64)
        self.count = int(count)
        self.min_b = np.asarray(min_b, dtype=np.int32)
        self.div_b = np.asarray(div_b, dtype=np.int32)
        self.leaf = float(np.float32(leaf))  # remove_duplicate takes a float leaf (point_cloud_helper.h:16)
        self.bounds = [int(b) for b in bounds]
        self.shard_counts = [int(c) for c in shard_counts]
        self.halo = float(halo)

    def face_x(self, col):
        """World x of the lower face of voxel column `col` (up to rounding; used only to place the halo)."""
        return float(self.centre[0]) + (col + int(self.min_b[0])) * self.leaf


def make_plan(ctx, chunks, world, leaf, halo):
    """chunks: callable returning an iterator of (device_ptr, n) pieces of the whole cloud in point order."""
    s3, cnt = np.zeros(3), 0
    mn, mx = np.full(3, np.inf), np.full(3, -np.inf)
    for ptr, n in chunks():
        s3, cnt = ctx.serial_sum((ptr, n, 48), s3, cnt)
        a, b = ctx.minmax((ptr, n, 48))
        mn, mx = np.minimum(mn, a), np.maximum(mx, b)
    centre = s3 / float(cnt)  # compute3DCentroid: sum / count (point_cloud_helper.h:227-229)
    leaf_d = float(np.float32(leaf))
    inv = 1.0 / leaf_d
    mn_c = mn + (-centre)                       # transformPointCloud: p + (-c)
    mx_c = np.maximum(mx + (-centre), DBL_MIN)  # getMinMax3D's max seed (point_cloud_helper.h:65)
    min_b = np.array([int(v * inv) for v in mn_c], dtype=np.int32)   # truncation, voxel_grid.h:835-840
    max_b = np.array([int(v * inv) for v in mx_c], dtype=np.int32)
    div_b = max_b - min_b + 1
    cpb = max(1, int(round(1.0 / leaf_d)))      # ~1 m of columns per histogram bin (CloudGrid's tile edge)
    nbins = int(div_b[0]) // cpb + 2
    hist = np.zeros(nbins, dtype=np.uint64)
    for ptr, n in chunks():
        ctx.column_histogram((ptr, n, 48), 0, leaf_d, centre[0], min_b[0], 0, cpb, hist)
    cum = np.cumsum(hist.astype(np.int64))
    total = int(cum[-1])
    bounds = [0]
    for r in range(1, world):
        b = int(np.searchsorted(cum, (total * r) // world, side="left")) + 1
        bounds.append(max(b * cpb, bounds[-1] + cpb))
    bounds.append(int(div_b[0]) + cpb * 2)
    counts = []
    for r in range(world):
        lo, hi = bounds[r] // cpb, min(bounds[r + 1] // cpb, nbins)
        counts.append(int(hist[lo:hi].sum()))
    return ShardPlan(centre, cnt, min_b, div_b, leaf, bounds, counts, halo)


class DistComm:
    """Neighbour exchange over torch.distributed (NCCL on GPUs, gloo in CPU tests)."""

    def __init__(self, dist, torch, rank, world, device, host_group=None):
        """host_group: optional gloo group for the message sizes, so that they travel host to host instead of
        through device memory (a device-to-host read would queue behind a bulk download on the copy engine)."""
        self.dist, self.torch, self.rank, self.world, self.device = dist, torch, rank, world, device
        self.host_group = host_group

    def exchange(self, send_l, send_r):
        """send_l / send_r: uint8 tensors for ranks rank-1 / rank+1 (None or empty allowed).
        Returns what those ranks sent to this one (empty tensors where there is no neighbour)."""
        t, d = self.torch, self.dist
        empty = t.empty(0, dtype=t.uint8, device=self.device)
        send_l = empty if send_l is None else send_l
        send_r = empty if send_r is None else send_r
        has_l, has_r = self.rank > 0, self.rank < self.world - 1
        sdev = "cpu" if self.host_group is not None else self.device
        grp = self.host_group
        sz_out = [t.tensor([send_l.numel()], dtype=t.int64, device=sdev),
                  t.tensor([send_r.numel()], dtype=t.int64, device=sdev)]
        sz_in = [t.zeros(1, dtype=t.int64, device=sdev), t.zeros(1, dtype=t.int64, device=sdev)]
        ops = []
        if has_l:
            ops += [d.P2POp(d.isend, sz_out[0], self.rank - 1, grp), d.P2POp(d.irecv, sz_in[0], self.rank - 1, grp)]
        if has_r:
            ops += [d.P2POp(d.isend, sz_out[1], self.rank + 1, grp), d.P2POp(d.irecv, sz_in[1], self.rank + 1, grp)]
        if ops:
            for w in d.batch_isend_irecv(ops):
                w.wait()
        nl, nr = int(sz_in[0].item()), int(sz_in[1].item())
        recv_l = t.empty(nl, dtype=t.uint8, device=self.device)
        recv_r = t.empty(nr, dtype=t.uint8, device=self.device)
        ops = []
        if has_l and send_l.numel():
            ops.append(d.P2POp(d.isend, send_l, self.rank - 1))
        if has_l and nl:
            ops.append(d.P2POp(d.irecv, recv_l, self.rank - 1))
        if has_r and send_r.numel():
            ops.append(d.P2POp(d.isend, send_r, self.rank + 1))
        if has_r and nr:
            ops.append(d.P2POp(d.irecv, recv_r, self.rank + 1))
        if ops:
            for w in d.batch_isend_irecv(ops):
                w.wait()
        return recv_l, recv_r


def run_lockstep(gens):
    """Drives the step generators of ALL ranks in one process (single-GPU emulation of the exchanges):
    rank r's send_l is delivered to rank r-1 as its recv_r, and so on.  Returns the generators' results."""
    world = len(gens)
    msgs = [next(g) for g in gens]
    results = [None] * world
    done = 0
    while done < world:
        recvs = []
        for r in range(world):
            from_l = msgs[r - 1][1] if r > 0 else None          # what r-1 sent to its right
            from_r = msgs[r + 1][0] if r < world - 1 else None  # what r+1 sent to its left
            recvs.append((from_l, from_r))
        nxt = []
        for r in range(world):
            try:
                nxt.append(gens[r].send(recvs[r]))
            except StopIteration as e:
                results[r] = e.value
                nxt.append((None, None))
                done += 1
        msgs = nxt
    return results


class ShardedPipeline:
    """remove_duplicate(leaf) -> index build -> kNN(k) + plane-fit features on this rank's shard."""

    CHUNK = 8_000_000  # points generated / scanned at a time while planning and cutting the cloud

    def __init__(self, ctx, rank, world, n_per_rank, seed=5, leaf=0.02, k=20, kind="city", comm=None, torch=None,
                 halo=1.0):
        self.ctx, self.rank, self.world = ctx, rank, world
        self.n_per_rank, self.seed, self.leaf, self.k, self.kind = int(n_per_rank), seed, leaf, k, kind
        self.comm, self.torch, self.halo = comm, torch, float(halo)
        self.n_total = self.n_per_rank * world
        self.plan = None
        self.n = 0
        self.cloud = self.work = self.props = self.kth = None  # device buffers (ptr attribute / torch tensors)
        self.cap = 0
        self.stats = {}
        self.last_m = 0
        self.bnd_cap, self.bnd_idx, self.bnd_xyz = 1 << 16, None, None  # boundary queries (grown on demand)

    # ---- buffers: torch tensors when torch is given (needed for the exchanges), library allocations otherwise ----
    def _alloc(self, nbytes):
        if self.torch is not None:
            return self.torch.empty(max(int(nbytes), 16), dtype=self.torch.uint8, device="cuda")
        return self.ctx.alloc(max(int(nbytes), 16))

    @staticmethod
    def _p(buf):
        return buf.data_ptr() if hasattr(buf, "data_ptr") else buf.ptr

    def _chunks(self, scratch):
        def it():
            first = 0
            while first < self.n_total:
                n = min(self.CHUNK, self.n_total - first)
                self.ctx.synth_cloud(self.kind, self.n_total, first=first, n=n, seed=self.seed, out=self._p(scratch))
                yield self._p(scratch), n
                first += n
        return it

    # ---- setup (untimed): generate the whole synthetic cloud chunk by chunk, keep this rank's slab ----
    def generate(self):
        ctx = self.ctx
        if self.world == 1:
            self.n = self.n_per_rank
            self.cloud = self._alloc(self.n * 48)
            ctx.synth_cloud(self.kind, self.n, seed=self.seed, out=self._p(self.cloud))
            self.cap = self.n
        else:
            scratch = self._alloc(min(self.CHUNK, self.n_total) * 48)
            chunks = self._chunks(scratch)
            self.plan = make_plan(ctx, chunks, self.world, self.leaf, self.halo)
            P = self.plan
            self.n = P.shard_counts[self.rank]
            self.cloud = self._alloc(self.n * 48)
            at = 0
            for ptr, n in chunks():
                at += ctx.select_columns((ptr, n, 48), 0, P.leaf, P.centre[0], P.min_b[0], P.bounds[self.rank],
                                         P.bounds[self.rank + 1], self._p(self.cloud) + at * 48)
            assert at == self.n, (at, self.n)
            del scratch
            self.cap = self.n + self.n // 4 + 1_000_000  # own + halos
            self.kth = self._alloc(self.cap * 4)
        self.work = self._alloc(self.cap * 48)
        self.props = self._alloc(self.cap * 48)
        ctx.synchronize()

    def halo_info(self):
        return self.stats or None

    # ---- one step ----
    def step_device(self):
        gen = self._step_gen(self._p(self.cloud))
        return self._run(gen) if self.world > 1 else self._drain(gen)

    def _run(self, gen):
        try:
            msg = next(gen)
            while True:
                msg = gen.send(self.comm.exchange(msg[0], msg[1]))
        except StopIteration as e:
            return e.value

    def _mark(self, name):
        """PCP_SHARD_TRACE=1: wall clock between the phases of a step (synchronises: a debugging aid)."""
        if not os.environ.get("PCP_SHARD_TRACE"):
            return
        import time
        self.torch.cuda.synchronize()
        now = time.perf_counter()
        if name == "start":
            self._trace = []
        else:
            self._trace.append("%s %.1f" % (name, (now - self._t_last) * 1e3))
        self._t_last = now
        if name == "end" and self.rank == 0:
            sys.stderr.write("[shard trace] " + ", ".join(self._trace) + "\n")

    def _step_gen(self, src_ptr):
        """Generator: yields (send_left, send_right) uint8 tensors at each exchange, receives (from_left,
        from_right); returns the number of featured points.  src_ptr: this rank's input cloud (device or pinned
        host memory)."""
        ctx, k = self.ctx, self.k
        wp = self._p(self.work)
        ctx.memcpy_async(wp, src_ptr, self.n * 48)  # remove_duplicate works in place: the input is kept
        if self.world == 1:
            m = ctx.remove_duplicate((wp, self.n, 48), self.leaf)
            ix = api.Index(ctx, (wp, m, 48))
            try:
                ix.features_knn(k, props=self._p(self.props), want_eig=False, want_lps=False)
            finally:
                ix.free()
            self.last_m = m
            return m
        t, P, H = self.torch, self.plan, self.halo
        self._mark("start")
        m = ctx.remove_duplicate_centred((wp, self.n, 48), P.leaf, P.centre, P.min_b, P.div_b)
        xlo, xhi = P.face_x(P.bounds[self.rank]), P.face_x(P.bounds[self.rank + 1])
        has_l, has_r = self.rank > 0, self.rank < self.world - 1
        # ---- exchange 1: voxelised points within H of the slab faces ----
        send_l = send_r = None
        if has_l:
            buf = t.empty(m * 48, dtype=t.uint8, device="cuda")
            c = ctx.select_range((wp, m, 48), 0, -math.inf, xlo + H, buf.data_ptr())
            send_l = buf[:c * 48]
        if has_r:
            buf = t.empty(m * 48, dtype=t.uint8, device="cuda")
            c = ctx.select_range((wp, m, 48), 0, xhi - H, math.inf, buf.data_ptr())
            send_r = buf[:c * 48]
        self._mark("dedup+select")
        recv_l, recv_r = yield (send_l, send_r)
        self._mark("exchange1")
        nl = 0 if recv_l is None else recv_l.numel() // 48
        nr = 0 if recv_r is None else recv_r.numel() // 48
        ntot = m + nl + nr
        if ntot > self.cap:
            raise RuntimeError("shard + halos (%d points) exceed the buffer (%d)" % (ntot, self.cap))
        if nl:
            self.work[m * 48:(m + nl) * 48].copy_(recv_l)
        if nr:
            self.work[(m + nl) * 48:ntot * 48].copy_(recv_r)
        ix = api.Index(ctx, (wp, ntot, 48))
        try:
            ix.features_knn(k, props=self._p(self.props), want_eig=False, want_lps=False, query_limit=m,
                            kth_d2=self._p(self.kth))
            self._mark("index+features")
            # ---- queries whose k-th neighbour may lie beyond the halo ----
            pts = self.work[:ntot * 48].view(t.float64).view(ntot, 6)
            while True:
                if self.bnd_idx is None or self.bnd_idx.numel() < self.bnd_cap:
                    self.bnd_idx = t.empty(self.bnd_cap, dtype=t.int32, device="cuda")
                    self.bnd_xyz = t.empty((self.bnd_cap, 3), dtype=t.float64, device="cuda")
                nF = ctx.select_boundary((wp, m, 48), 0, self._p(self.kth), (xlo - H) if has_l else None,
                                         (xhi + H) if has_r else None, self.bnd_cap, self.bnd_idx.data_ptr(),
                                         self.bnd_xyz.data_ptr())
                if nF >= 0:
                    break
                self.bnd_cap = -nF + (-nF) // 4
            F = self.bnd_idx[:nF]
            qF = self.bnd_xyz[:nF]
            check = os.environ.get("PCP_SHARD_CHECK")  # debugging aid: validate every index before it is used
            if check and nF:
                assert int(F.min()) >= 0 and int(F.max()) < m, ("F", self.rank, int(F.min()), int(F.max()), m, nF)
            # each flagged query goes only to the neighbour(s) whose halo it outreaches (a query sent to the far
            # side would be a whole slab away from that rank's points: correct, but a search over everything)
            kthF = self.kth[:m * 4].view(t.float32)[F.long()].contiguous()
            side = []
            for has, lo, hi in ((has_l, xlo - H, None), (has_r, None, xhi + H)):
                if not has or nF == 0:
                    side.append(None)
                    continue
                sub = t.empty(nF, dtype=t.int32, device="cuda")
                sxyz = t.empty((nF, 3), dtype=t.float64, device="cuda")
                c = ctx.select_boundary((qF.data_ptr(), nF, 24), 0, kthF.data_ptr(), lo, hi, nF, sub.data_ptr(),
                                        sxyz.data_ptr())
                side.append((sub[:c], sxyz[:c]))
                if check and c:
                    assert 0 <= c <= nF and int(sub[:c].min()) >= 0 and int(sub[:c].max()) < nF, \
                        ("side", self.rank, c, nF, int(sub[:c].min()), int(sub[:c].max()))
            # ---- exchange 2: the flagged queries ----
            self._mark("boundary select")
            q_l, q_r = yield tuple(None if sd is None else sd[1].view(t.uint8).view(-1) for sd in side)
            self._mark("exchange2")
            answers = []
            for q in (q_l, q_r):
                if q is None or q.numel() == 0:
                    answers.append(None)
                    continue
                nq = q.numel() // 24
                qq = q.contiguous()
                idx = t.empty((nq, k), dtype=t.int32, device="cuda")
                d2 = t.empty((nq, k), dtype=t.float64, device="cuda")
                ix.knn_device(qq.data_ptr(), nq, 24, k, idx.data_ptr(), d2.data_ptr())
                if check:
                    assert int(idx.max()) < ntot and int(idx.min()) >= -1, ("answer idx", self.rank, int(idx.min()), int(idx.max()), ntot, nq)
                xyz = pts[idx.view(-1).long().clamp(min=0), :3].contiguous()  # padded entries carry d2 = +inf
                answers.append(t.cat([xyz.view(t.uint8).view(-1), d2.view(t.uint8).view(-1)]))
            # ---- exchange 3: candidate lists back to the owners ----
            self._mark("answers")
            a_l, a_r = yield (answers[0], answers[1])
            self._mark("exchange3")
            if nF:
                idx = t.empty((nF, k), dtype=t.int32, device="cuda")
                d2o = t.empty((nF, k), dtype=t.float64, device="cuda")
                ix.knn_device(qF.data_ptr(), nF, 24, k, idx.data_ptr(), d2o.data_ptr())
                if check:
                    assert int(idx.max()) < ntot and int(idx.min()) >= -1, ("own idx", self.rank, int(idx.min()), int(idx.max()), ntot, nF)
                xyz_lists = [pts[idx.view(-1).long().clamp(min=0), :3].contiguous().view(-1)]
                d2_lists = [d2o.view(-1)]
                for a, sd in ((a_l, side[0]), (a_r, side[1])):
                    if a is None or a.numel() == 0 or sd is None:
                        continue
                    rows = sd[0].long()
                    c = rows.numel()
                    assert a.numel() == c * k * 32, (a.numel(), c, k)
                    # queries that were not sent to this side get an empty list (d2 = +inf)
                    xyz_full = t.zeros((nF, k, 3), dtype=t.float64, device="cuda")
                    d2_full = t.full((nF, k), math.inf, dtype=t.float64, device="cuda")
                    xyz_full[rows] = a[:c * k * 24].contiguous().view(t.float64).view(c, k, 3)
                    d2_full[rows] = a[c * k * 24:].contiguous().view(t.float64).view(c, k)
                    xyz_lists.append(xyz_full.view(-1))
                    d2_lists.append(d2_full.view(-1))
                nlists = len(xyz_lists)
                xyz_all, d2_all, pid = t.cat(xyz_lists), t.cat(d2_lists), F
                ctx.features_merge(nF, k, nlists, xyz_all.data_ptr(), d2_all.data_ptr(), pid.data_ptr(),
                                   self._p(self.props))
                del xyz_all, d2_all, pid
        finally:
            ix.free()
        self._mark("merge")
        self._mark("end")
        self.stats = {"halo_m": H, "own": int(m), "halo_points": int(nl + nr), "boundary_queries": nF,
                      "sent": [0 if sd is None else int(sd[0].numel()) for sd in side]}
        self.last_m = m
        return m

    # ---- end to end: host buffers through the C ABI ----
    def time_e2e(self, steps, barrier, sync, keep=None):
        """Times `steps` passes whose input cloud starts in pinned HOST memory and whose LAS_POINT_PROPERTY
        records end in pinned HOST memory, every byte moved through the C ABI inside the timed region.  The
        passes are pipelined over the library's copy lanes (pcp_copy_async): while pass i is processed, the
        cloud of pass i+1 is uploaded and the records of pass i-1 are downloaded, each pass with its own device
        and host buffers.  Wall clock around all of it (fill and drain included), barrier + device
        synchronisation on both sides.  Returns (ms_per_step, h2d_bytes, d2h_bytes, featured per step).
        keep: optional list that receives a copy of the host records of the last two passes (tests)."""
        import time
        ctx = self.ctx
        nb = self.n * 48
        host_in, pin_in = ctx.pinned(nb)
        pins = [ctx.pinned(self.cap * 48) for _ in range(2)]
        dev_in = [self._alloc(nb) for _ in range(2)]
        props = [self.props, self._alloc(self.cap * 48)]
        ctx.memcpy(pin_in, self._p(self.cloud), nb)
        UP, DOWN = 0, (1, 2)

        def run(nsteps):
            feat = d2h = 0
            ctx.copy_async(UP, self._p(dev_in[0]), pin_in, nb)
            for i in range(nsteps):
                b = i & 1
                ctx.copy_fence(UP)            # pass i reads dev_in[b]: wait for its upload
                ctx.copy_fence(DOWN[b])       # ... and writes props[b]: wait for the download of pass i-2
                if i + 1 < nsteps:            # dev_in[1-b] was last read by pass i-1, already submitted
                    ctx.copy_async(UP, self._p(dev_in[1 - b]), pin_in, nb)
                self.props = props[b]
                gen = self._step_gen(self._p(dev_in[b]))
                m = self._run(gen) if self.world > 1 else self._drain(gen)
                ctx.copy_async(DOWN[b], pins[b][1], self._p(props[b]), m * 48)
                feat += m
                d2h = m * 48
            ctx.copy_wait(DOWN[0])
            ctx.copy_wait(DOWN[1])
            return feat, d2h

        try:
            run(2)  # warms the arena, the lanes and both buffer sets
            sync()
            barrier()
            t0 = time.perf_counter()
            feat, d2h = run(steps)
            sync()
            barrier()
            ms = (time.perf_counter() - t0) * 1e3 / steps
            if keep is not None:
                keep.extend(pins[b][0][:d2h].copy() for b in ((steps - 1) & 1, steps & 1)[:min(steps, 2)])
        finally:
            self.props = props[0]
            ctx.synchronize()
            ctx.free_pinned(pin_in)
            for _, p in pins:
                ctx.free_pinned(p)
            for a in dev_in + [props[1]]:
                if hasattr(a, "free"):
                    a.free()
        return ms, nb, d2h, feat / steps

    @staticmethod
    def _drain(gen):
        try:
            next(gen)
        except StopIteration as e:
            return e.value
        raise RuntimeError("a single-rank step must not exchange")

    def close(self):
        for a in (self.cloud, self.work, self.props, self.kth):
            if a is not None and hasattr(a, "free"):
                a.free()
        self.cloud = self.work = self.props = self.kth = None
