"""The voxel-grid -> kNN -> feature pipeline over one or several GPUs: the host-side driver that bench.py
and the multi-GPU tests run, one process per GPU.

Input: every rank holds a CONTIGUOUS RANGE of the cloud in point order (rank r: points [r n/N, (r+1) n/N)), as packed
32-byte records (x, y, z, rgba, stamp_id: the binary PCD payload record) -- however the points are spread in space.
One step, everything inside:
  plan          per-rank box / count / tree sums (one all-gather), the serial-order centroid of the WHOLE cloud
                (compute3DCentroid, point_cloud_helper.h:192-230: every rank prepares its range in parallel, the exact
                running sum then walks from rank to rank -- tens of microseconds each), the grid box of the centred
                cloud, a histogram of voxel columns along x in ~1 m bins (CloudGrid's tile edge, cloud_grid.cpp:34-78;
                one all-reduce), slabs of equal point count cut on column boundaries;
  redistribute  a stable partition by destination slab and one all-to-all over NVLink: every voxel's points end up
                on one rank, in point order;
  shard         remove_duplicate centred on the whole cloud's centroid in the whole cloud's box (cells, centroids
                and order equal the single-device result, no exchange); halo exchange with both neighbours (voxelised
                points within `halo` of the slab faces); index + kNN + plane fit on own + halo points; the queries
                whose k-th neighbour may lie beyond the halo are answered exactly by the neighbour and merged.
All compute is libpcp_cuda.so; torch is device buffers, streams and torch.distributed only.

The step is a generator that yields its communication requests, so the same code runs under torch.distributed
(DistComm.run: NCCL, or gloo in CPU tests) and, for single-GPU parity tests, with all ranks in one process
(run_lockstep).  Requests: ("allgather", t) ("allreduce", t) ("bcast", (t, src)) ("chain_recv", t) ("chain_send", t)
("alltoall", (buf, byte_counts, out)) ("exchange", (left, right)) ("gather", (t, nbytes)).
"""
import math
import os
import sys

import numpy as np

from . import api

DBL_MIN = 2.2250738585072014e-308
REC = 32  # bytes of a packed cloud record


class ShardPlan:
    """What every rank must agree on: centre, grid box, slab boundaries (voxel columns along x)."""

    def __init__(self, centre, count, min_b, div_b, leaf, bounds, halo, axis=0):
        self.axis = int(axis)  # the slabs are cut along the longer horizontal extent of the cloud (0 = x, 1 = y)
        self.centre = np.asarray(centre, dtype=np.float64)
        self.count = int(count)
        self.min_b = np.asarray(min_b, dtype=np.int32)
        self.div_b = np.asarray(div_b, dtype=np.int32)
        self.leaf = float(np.float32(leaf))  # remove_duplicate takes a float leaf (point_cloud_helper.h:16)
        self.bounds = [int(b) for b in bounds]
        self.halo = float(halo)

    def face_x(self, col):
        """World coordinate (along the cut axis) of the lower face of voxel column `col` (up to rounding; used only to
        place the halo)."""
        return float(self.centre[self.axis]) + (col + int(self.min_b[self.axis])) * self.leaf


def grid_box(mn, mx, centre, leaf):
    """min_b / div_b of the cloud translated by -centre (voxel_grid.h:835-847 after point_cloud_helper.cpp:52-53)."""
    leaf_d = float(np.float32(leaf))
    inv = 1.0 / leaf_d
    mn_c = np.asarray(mn) + (-np.asarray(centre))                       # transformPointCloud: p + (-c)
    mx_c = np.maximum(np.asarray(mx) + (-np.asarray(centre)), DBL_MIN)  # getMinMax3D's max seed (point_cloud_helper.h:65)
    min_b = np.array([int(v * inv) for v in mn_c], dtype=np.int32)      # truncation, voxel_grid.h:835-840
    max_b = np.array([int(v * inv) for v in mx_c], dtype=np.int32)
    return min_b, max_b - min_b + 1


def cut_slabs(hist, world, cpb, div_x, leaf, halo):
    """Slab boundaries (voxel columns) of equal point count on histogram-bin boundaries.  Every slab is at least `halo`
    wide: a rank receives halo points from its direct neighbours only, so a narrower slab would leave the points of
    rank r +- 2 inside [face - halo, face] unseen (and the queries that reach them unflagged)."""
    leaf_d = float(np.float32(leaf))
    min_bins = max(1, int(math.ceil(halo / (cpb * leaf_d) - 1e-9)))
    nbins = len(hist)
    cum = np.cumsum(np.asarray(hist, dtype=np.int64))
    total = int(cum[-1])
    last_bin = int(div_x) // cpb + 1
    if world > 1 and world * min_bins > last_bin + 1:
        raise ValueError("cloud too narrow along x for %d slabs at least %.3g m (halo) wide" % (world, halo))
    b = [0]
    for r in range(1, world):
        want = int(np.searchsorted(cum, (total * r) // world, side="left")) + 1
        lo = b[-1] + min_bins                          # this slab at least min_bins wide ...
        hi = last_bin + 1 - (world - r) * min_bins     # ... and room for the ones after it
        b.append(min(max(want, lo), max(hi, lo)))
    b.append(nbins)
    bounds = [x * cpb for x in b]
    for r in range(world):
        if world > 1 and (bounds[r + 1] - bounds[r]) * leaf_d < halo - 1e-9:
            raise ValueError("slab %d is %.3g m wide, less than the halo %.3g m" % (r, (bounds[r + 1] - bounds[r]) * leaf_d, halo))
    return bounds


# ---- communication ---------------------------------------------------------------------------------------------
class DistComm:
    """The step's requests over torch.distributed: NCCL on GPUs (device tensors), gloo in the CPU tests.  Message sizes
    travel in band (a one-word message before the payload), read back through `peek` -- on a GPU the library's mapped
    read-back, which does not queue behind a bulk download on the copy engine."""

    def __init__(self, dist, torch, rank, world, device, peek=None, poke=None):
        self.dist, self.torch, self.rank, self.world, self.device = dist, torch, rank, world, device
        self.peek = peek or (lambda t: t.cpu().numpy())
        # poke(device tensor, numpy array): small host values into device memory.  On a GPU the library's mapped write:
        # a plain copy takes a place in the copy engine's queue BEHIND the bulk upload of the next pass, and every
        # size word of every exchange then waits for that upload -- the passes and their uploads ran one after the other
        self.poke = poke or (lambda ten, arr: ten.copy_(torch.from_numpy(np.ascontiguousarray(arr)).view(ten.dtype).reshape(ten.shape)))
        self._sz = None
        self._cnt = None

    # -- small helpers --
    def _p2p(self, ops):
        if ops:
            for w in self.dist.batch_isend_irecv(ops):
                w.wait()

    def _sizes(self, pairs):
        """pairs: list of (peer, nbytes to send).  Returns the byte counts the peers send to this rank."""
        t, d = self.torch, self.dist
        if self._sz is None:
            self._sz = (t.zeros(2 * self.world, dtype=t.int64, device=self.device), t.zeros(2 * self.world, dtype=t.int64, device=self.device))
        out, inn = self._sz
        host = np.zeros(2 * self.world, dtype=np.int64)
        for j, (peer, nb) in enumerate(pairs):
            host[j] = nb
        self.poke(out, host)
        ops = []
        for j, (peer, nb) in enumerate(pairs):
            ops += [d.P2POp(d.isend, out[j:j + 1], peer), d.P2POp(d.irecv, inn[j:j + 1], peer)]
        self._p2p(ops)
        got = self.peek(inn[:max(len(pairs), 1)])
        return [int(got[j]) for j in range(len(pairs))]

    # -- requests --
    def handle(self, req):
        t, d = self.torch, self.dist
        kind, arg = req
        if kind == "allgather":
            out = t.empty((self.world,) + tuple(arg.shape), dtype=arg.dtype, device=arg.device)
            d.all_gather_into_tensor(out, arg.contiguous()) if arg.is_cuda else d.all_gather(list(out.unbind(0)), arg.contiguous())
            return out
        if kind == "allreduce":
            d.all_reduce(arg)
            return arg
        if kind == "bcast":
            ten, src = arg
            d.broadcast(ten, src)
            return ten
        if kind == "chain_recv":
            if self.rank > 0:
                d.recv(arg, self.rank - 1)
            else:
                arg.zero_()
            return arg
        if kind == "chain_send":
            if self.rank < self.world - 1:
                d.send(arg, self.rank + 1)
            return None
        if kind == "exchange":
            left, right = arg
            peers = []
            if self.rank > 0:
                peers.append((self.rank - 1, left))
            if self.rank < self.world - 1:
                peers.append((self.rank + 1, right))
            sizes = self._sizes([(p, 0 if s is None else s.numel()) for p, s in peers])
            ops, recv = [], {}
            for (p, s), nb in zip(peers, sizes):
                if s is not None and s.numel():
                    ops.append(d.P2POp(d.isend, s, p))
                recv[p] = t.empty(nb, dtype=t.uint8, device=self.device)
                if nb:
                    ops.append(d.P2POp(d.irecv, recv[p], p))
            self._p2p(ops)
            return recv.get(self.rank - 1), recv.get(self.rank + 1)
        if kind == "alltoall":
            buf, counts, out = arg  # counts[q]: bytes for rank q, consecutive in buf; out: receive buffer (uint8)
            if self._cnt is None:
                self._cnt = t.zeros(self.world, dtype=t.int64, device=self.device)
            cnt = self._cnt
            self.poke(cnt, np.asarray(counts, dtype=np.int64))
            allc = self.handle(("allgather", cnt))
            m = np.asarray(self.peek(allc)).reshape(self.world, self.world)  # m[src][dst]
            recv_counts = [int(m[q][self.rank]) for q in range(self.world)]
            total = sum(recv_counts)
            if total > out.numel():
                raise RuntimeError("all-to-all: %d bytes arrive, room for %d" % (total, out.numel()))
            ops, so, ro = [], 0, 0
            for q in range(self.world):
                sb, rb = int(counts[q]), recv_counts[q]
                if q == self.rank:
                    out[ro:ro + rb].copy_(buf[so:so + sb])
                else:
                    if sb:
                        ops.append(d.P2POp(d.isend, buf[so:so + sb], q))
                    if rb:
                        ops.append(d.P2POp(d.irecv, out[ro:ro + rb], q))
                so += sb
                ro += rb
            self._p2p(ops)
            return total
        if kind == "gather":
            ten, nbytes = arg  # to rank 0: returns (uint8 tensor, byte offsets per rank) there, None elsewhere
            if self._cnt is None:
                self._cnt = t.zeros(self.world, dtype=t.int64, device=self.device)
            cnt = self._cnt[:1]
            self.poke(cnt, np.asarray([int(nbytes)], dtype=np.int64))
            allc = [int(v) for v in np.asarray(self.peek(self.handle(("allgather", cnt)))).reshape(-1)]
            if self.rank == 0:
                off = np.concatenate([[0], np.cumsum(allc)]).astype(np.int64)
                out = t.empty(int(off[-1]), dtype=t.uint8, device=self.device)
                out[:allc[0]].copy_(ten[:allc[0]])
                self._p2p([d.P2POp(d.irecv, out[int(off[q]):int(off[q + 1])], q) for q in range(1, self.world) if allc[q]])
                return out, off
            if nbytes:
                self._p2p([d.P2POp(d.isend, ten[:int(nbytes)], 0)])
            return None
        raise ValueError("unknown request %r" % (kind,))

    def run(self, gen):
        try:
            req = next(gen)
            while True:
                req = gen.send(self.handle(req))
        except StopIteration as e:
            return e.value


def run_lockstep(gens, torch=None):
    """Drives the step generators of ALL ranks in one process (single-GPU emulation of the collectives).  A request is
    answered when it can be: collectives once every live rank has posted the same one, the chain as the carry arrives.
    Returns the generators' results."""
    world = len(gens)
    req, res, done = [None] * world, [None] * world, [False] * world
    box = {}

    def advance(r, value, first=False):
        try:
            req[r] = next(gens[r]) if first else gens[r].send(value)
        except StopIteration as e:
            res[r], done[r], req[r] = e.value, True, None

    for r in range(world):
        advance(r, None, first=True)
    while not all(done):
        progressed = False
        for r in range(world):
            while not done[r] and req[r][0] in ("chain_recv", "chain_send"):
                kind, ten = req[r]
                if kind == "chain_send":
                    box[r + 1] = ten.clone()
                    advance(r, None)
                elif r == 0 or r in box:
                    if r == 0:
                        ten.zero_()
                    else:
                        ten.copy_(box.pop(r))
                    advance(r, ten)
                else:
                    break
                progressed = True
        live = [r for r in range(world) if not done[r]]
        kinds = {req[r][0] for r in live}
        if live and len(kinds) == 1 and len(live) == world and not (kinds & {"chain_recv", "chain_send"}):
            kind = kinds.pop()
            args = [req[r][1] for r in range(world)]
            if kind == "allgather":
                stack = args[0].new_empty((world,) + tuple(args[0].shape))
                for r in range(world):
                    stack[r].copy_(args[r])
                out = [stack.clone() for _ in range(world)]
            elif kind == "allreduce":
                tot = args[0].clone()
                for r in range(1, world):
                    tot += args[r]
                for r in range(world):
                    args[r].copy_(tot)
                out = args
            elif kind == "bcast":
                src = args[0][1]
                for r in range(world):
                    args[r][0].copy_(args[src][0])
                out = [a[0] for a in args]
            elif kind == "exchange":
                out = []
                for r in range(world):
                    fl = args[r - 1][1] if r > 0 else None
                    fr = args[r + 1][0] if r < world - 1 else None
                    out.append((None if fl is None else fl.clone(), None if fr is None else fr.clone()))
            elif kind == "alltoall":
                out = []
                for r in range(world):
                    ro = 0
                    for q in range(world):
                        buf, counts, _ = args[q]
                        so = int(sum(int(c) for c in counts[:r]))
                        nb = int(counts[r])
                        args[r][2][ro:ro + nb].copy_(buf[so:so + nb])
                        ro += nb
                    out.append(ro)
            elif kind == "gather":
                sizes = [int(a[1]) for a in args]
                off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
                cat = args[0][0].new_empty(int(off[-1]))
                for r in range(world):
                    cat[int(off[r]):int(off[r + 1])].copy_(args[r][0][:sizes[r]])
                out = [(cat, off)] + [None] * (world - 1)
            else:
                raise ValueError("unknown request %r" % (kind,))
            for r in range(world):
                advance(r, out[r])
            progressed = True
        if not progressed:
            raise RuntimeError("lock-step deadlock: %r" % ([None if q is None else q[0] for q in req],))
    return res


# ---- the pipeline ----------------------------------------------------------------------------------------------
class ShardedPipeline:
    """remove_duplicate(leaf) -> index build -> kNN(k) + plane-fit features of one cloud over `world` GPUs."""

    CHUNK = 8_000_000  # points generated at a time (setup)

    def __init__(self, ctx, rank, world, n_total, seed=5, leaf=0.02, k=20, kind="city", comm=None, torch=None, halo=2.0):
        if torch is None:
            raise ValueError("ShardedPipeline needs torch for its device buffers")
        self.ctx, self.rank, self.world = ctx, rank, world
        self.n_total, self.seed, self.leaf, self.k, self.kind = int(n_total), seed, leaf, k, kind
        self.comm, self.torch, self.halo = comm, torch, float(halo)
        # this rank's range of the cloud in point order
        self.first = (self.n_total * rank) // world
        self.n = (self.n_total * (rank + 1)) // world - self.first
        self.plan = None
        self.src = None      # packed input records of this rank's range (device)
        self.cloud = None    # this rank's slab after the redistribution (world > 1)
        self.part = None     # partition buffer (world > 1)
        self.work = None     # voxelised own points, then the halo points
        self.props = None    # LAS_POINT_PROPERTY records of the own points
        self.packed = None   # the same as 24-byte records (e2e)
        self.kth = None
        self.cap = 0
        self.stats = {}
        self.last_m = 0
        self.bnd_cap, self.bnd_idx, self.bnd_xyz = 1 << 16, None, None  # boundary queries (grown on demand)
        self._small = None
        # the library's kernels and torch's ops (exchanges, gathers) touch the same buffers: one stream for both
        ctx.set_stream(torch.cuda.current_stream().cuda_stream)

    # ---- buffers ----
    def _alloc(self, nbytes):
        return self.torch.empty(max(int(nbytes), 16), dtype=self.torch.uint8, device="cuda")

    @staticmethod
    def _p(buf):
        return buf.data_ptr() if hasattr(buf, "data_ptr") else buf.ptr

    def _poke(self, ten, arr):
        """A small host array into a device tensor through the library's mapped write (pcp_poke)."""
        if ten.is_cuda:
            self.ctx.poke(ten.data_ptr(), np.ascontiguousarray(arr))
        else:
            ten.copy_(self.torch.from_numpy(np.ascontiguousarray(arr)).view(ten.dtype).reshape(ten.shape))

    def _peek(self, ten):
        """A small device tensor on the host, through the library's mapped read-back (<= 64 KiB), else a plain copy."""
        nb = ten.numel() * ten.element_size()
        dt = {self.torch.int64: np.int64, self.torch.float64: np.float64, self.torch.int32: np.int32,
              self.torch.uint8: np.uint8}[ten.dtype]
        if nb <= 65536 and ten.is_cuda:
            return self.ctx.peek(ten.contiguous().data_ptr(), dt, ten.numel()).reshape(tuple(ten.shape))
        return ten.cpu().numpy()

    # ---- setup (untimed): this rank's range of the synthetic cloud, as packed records in HBM ----
    def generate(self):
        ctx, t = self.ctx, self.torch
        self.src = self._alloc(self.n * REC)
        scratch = self._alloc(min(self.CHUNK, max(self.n, 1)) * 48)
        at = 0
        while at < self.n:
            c = min(self.CHUNK, self.n - at)
            ctx.synth_cloud(self.kind, self.n_total, first=self.first + at, n=c, seed=self.seed, out=scratch.data_ptr())
            ctx.cloud_pack(scratch.data_ptr(), c, self.src.data_ptr() + at * REC)
            at += c
        ctx.synchronize()
        del scratch
        self._buffers()

    def set_input(self, packed_host):
        """This rank's range from a host array of PACKED_POINT_DTYPE records (tests)."""
        assert len(packed_host) == self.n
        self.src = self._alloc(self.n * REC)
        self.ctx.memcpy(self.src.data_ptr(), packed_host.ctypes.data, self.n * REC)
        self._buffers()

    def _buffers(self):
        if self.world == 1:
            self.cap = self.n
        else:
            # a slab holds ~ n_total / world points (the cut balances them up to one histogram bin), plus the halos
            self.cap = int(self.n_total / self.world * 1.25) + 2_000_000
            self.cloud = self._alloc(self.cap * REC)
            self.part = self._alloc(self.n * REC)
            self.kth = self._alloc(self.cap * 4)
            self._small = {"stats": self.torch.zeros(10, dtype=self.torch.float64, device="cuda"),
                           "carry": self.torch.zeros(3, dtype=self.torch.float64, device="cuda"),
                           "hist": self.torch.zeros(1 << 16, dtype=self.torch.int64, device="cuda")}
        self.work = self._alloc(self.cap * REC)
        self.props = self._alloc(self.cap * 48)
        self.ctx.synchronize()

    def halo_info(self):
        return self.stats or None

    # ---- one step ----
    def step_device(self, src_ptr=None):
        gen = self._step_gen(self._p(self.src) if src_ptr is None else src_ptr)
        return self.comm.run(gen) if self.world > 1 else self._drain(gen)

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
            self._trace.append("%s %.2f" % (name, (now - self._t_last) * 1e3))
        self._t_last = now
        if name == "end" and self.rank == 0:
            sys.stderr.write("[shard trace] " + ", ".join(self._trace) + "\n")

    def _plan_gen(self, src_ptr):
        """The plan phase (collectives) -> ShardPlan.  Every rank computes the same plan from the same reduced data."""
        ctx, t, S = self.ctx, self.torch, self._small
        cloud = (src_ptr, self.n, REC)
        ctx.cloud_stats(cloud, out=S["stats"].data_ptr())
        allst = self._peek((yield ("allgather", S["stats"]))).reshape(self.world, 10)
        pred = allst[:self.rank, 6:9].sum(axis=0) if self.rank else np.zeros(3)
        h = ctx.serial_sum_begin(cloud, pred)           # parallel on every rank ...
        yield ("chain_recv", S["carry"])
        ctx.serial_sum_end(h, S["carry"].data_ptr())    # ... the exact running sum walks from rank to rank
        yield ("chain_send", S["carry"])
        yield ("bcast", (S["carry"], self.world - 1))
        sums = self._peek(S["carry"])
        count = int(allst[:, 9].sum())
        centre = sums / float(count)                    # compute3DCentroid: sum / count (point_cloud_helper.h:227-229)
        mn, mx = allst[:, 0:3].min(axis=0), allst[:, 3:6].max(axis=0)
        leaf = self.leaf or 0.02  # without a voxel grid the columns only serve the cut
        min_b, div_b = grid_box(mn, mx, centre, leaf)
        # slabs along the longer horizontal side: a 2 km x 30 m strip scanned along y is cut along y, not into eight
        # slabs 4 m thick and 2 km long
        ax = 0 if (mx[0] - mn[0]) >= (mx[1] - mn[1]) else 1
        leaf_d = float(np.float32(leaf))
        cpb = max(1, int(round(1.0 / leaf_d)))          # ~1 m of columns per histogram bin (CloudGrid's tile edge)
        nbins = int(div_b[ax]) // cpb + 2
        if nbins > S["hist"].numel():
            raise RuntimeError("cloud spans %d one-metre columns along x, more than the histogram holds" % nbins)
        hist = S["hist"][:nbins]
        hist.zero_()
        ctx.column_histogram(cloud, ax, leaf_d, centre[ax], min_b[ax], 0, cpb, (hist.data_ptr(), nbins))
        yield ("allreduce", hist)
        bounds = cut_slabs(self._peek(hist), self.world, cpb, int(div_b[ax]), leaf, self.halo)
        return ShardPlan(centre, count, min_b, div_b, leaf, bounds, self.halo, ax)

    def _step_gen(self, src_ptr):
        """Generator: yields communication requests (module docstring); returns the number of featured points.
        src_ptr: this rank's range of the cloud, packed records in device memory."""
        ctx, k, t = self.ctx, self.k, self.torch
        wp = self._p(self.work)
        if self.world == 1:
            if self.leaf:
                m = ctx.remove_duplicate_to((src_ptr, self.n, REC), self.leaf, wp)
            else:  # no voxel grid in this workload: features of the cloud as it is
                m = self.n
                ctx.memcpy_async(wp, src_ptr, m * REC)
            ix = api.Index(ctx, (wp, m, REC))
            try:
                ix.features_knn(k, props=self._p(self.props), want_eig=False, want_lps=False)
            finally:
                ix.free()
            self.last_m = m
            return m
        self._mark("start")
        P = self.plan = yield from self._plan_gen(src_ptr)
        H = self.halo
        self._mark("plan")
        # ---- redistribute: every voxel's points to the rank that owns its column, in point order ----
        ax = P.axis
        counts = ctx.partition_columns((src_ptr, self.n, REC), ax, P.leaf, P.centre[ax], P.min_b[ax], P.bounds, self._p(self.part))
        nbytes = yield ("alltoall", (self.part, [int(c) * REC for c in counts], self.cloud))
        n_own = nbytes // REC
        self._mark("redistribute")
        if self.leaf:
            m = ctx.remove_duplicate_to((self._p(self.cloud), n_own, REC), P.leaf, wp, P.centre, P.min_b, P.div_b)
        else:
            m = n_own
            ctx.memcpy_async(wp, self._p(self.cloud), m * REC)
        xlo, xhi = P.face_x(P.bounds[self.rank]), P.face_x(P.bounds[self.rank + 1])
        has_l, has_r = self.rank > 0, self.rank < self.world - 1
        # ---- exchange 1: voxelised points within H of the slab faces (staged in the free part of `cloud`) ----
        send_l = send_r = None
        room = self.cap - n_own
        stage = self.cloud[n_own * REC:]
        cl = cr = 0
        if has_l:
            cl = ctx.select_range((wp, m, REC), ax, -math.inf, xlo + H, stage.data_ptr(), cap=room)
            send_l = stage[:cl * REC]
        if has_r:
            cr = ctx.select_range((wp, m, REC), ax, xhi - H, math.inf, stage.data_ptr() + cl * REC, cap=room - cl)
            send_r = stage[cl * REC:(cl + cr) * REC]
        self._mark("dedup+select")
        recv_l, recv_r = yield ("exchange", (send_l, send_r))
        self._mark("exchange1")
        nl = 0 if recv_l is None else recv_l.numel() // REC
        nr = 0 if recv_r is None else recv_r.numel() // REC
        ntot = m + nl + nr
        if ntot > self.cap:
            raise RuntimeError("shard + halos (%d points) exceed the buffer (%d)" % (ntot, self.cap))
        if nl:
            self.work[m * REC:(m + nl) * REC].copy_(recv_l)
        if nr:
            self.work[(m + nl) * REC:ntot * REC].copy_(recv_r)
        ix = api.Index(ctx, (wp, ntot, REC))
        nF, side = 0, [None, None]
        try:
            ix.features_knn(k, props=self._p(self.props), want_eig=False, want_lps=False, query_limit=m,
                            kth_d2=self._p(self.kth))
            self._mark("index+features")
            # ---- queries whose k-th neighbour may lie beyond the halo ----
            pts = self.work[:ntot * REC].view(t.float64).view(ntot, 4)
            while True:
                if self.bnd_idx is None or self.bnd_idx.numel() < self.bnd_cap:
                    self.bnd_idx = t.empty(self.bnd_cap, dtype=t.int32, device="cuda")
                    self.bnd_xyz = t.empty((self.bnd_cap, 3), dtype=t.float64, device="cuda")
                    self.bnd_sub = [t.empty(self.bnd_cap, dtype=t.int32, device="cuda") for _ in range(2)]
                    self.bnd_sxyz = [t.empty((self.bnd_cap, 3), dtype=t.float64, device="cuda") for _ in range(2)]
                nF = ctx.select_boundary((wp, m, REC), ax, self._p(self.kth), (xlo - H) if has_l else None,
                                         (xhi + H) if has_r else None, self.bnd_cap, self.bnd_idx.data_ptr(),
                                         self.bnd_xyz.data_ptr())
                if nF >= 0:
                    break
                self.bnd_cap = -nF + (-nF) // 4
            F = self.bnd_idx[:nF]
            qF = self.bnd_xyz[:nF]
            # each flagged query goes only to the neighbour(s) whose halo it outreaches (a query sent to the far side
            # would be a whole slab away from that rank's points: correct, but a search over everything)
            kthF = self.kth[:m * 4].view(t.float32)[F.long()].contiguous() if nF else None
            for s, (has, lo, hi) in enumerate(((has_l, xlo - H, None), (has_r, None, xhi + H))):
                if not has or nF == 0:
                    continue
                c = ctx.select_boundary((qF.data_ptr(), nF, 24), ax, kthF.data_ptr(), lo, hi, nF, self.bnd_sub[s].data_ptr(),
                                        self.bnd_sxyz[s].data_ptr())
                side[s] = (self.bnd_sub[s][:c], self.bnd_sxyz[s][:c])
            # ---- exchange 2: the flagged queries ----
            self._mark("boundary select")
            q_l, q_r = yield ("exchange", tuple(None if sd is None else sd[1].view(t.uint8).view(-1) for sd in side))
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
                xyz = pts[idx.view(-1).long().clamp(min=0), :3].contiguous()  # padded entries carry d2 = +inf
                answers.append(t.cat([xyz.view(t.uint8).view(-1), d2.view(t.uint8).view(-1)]))
            # ---- exchange 3: candidate lists back to the owners ----
            self._mark("answers")
            a_l, a_r = yield ("exchange", (answers[0], answers[1]))
            self._mark("exchange3")
            if nF:
                idx = t.empty((nF, k), dtype=t.int32, device="cuda")
                d2o = t.empty((nF, k), dtype=t.float64, device="cuda")
                ix.knn_device(qF.data_ptr(), nF, 24, k, idx.data_ptr(), d2o.data_ptr())
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
                xyz_all, d2_all = t.cat(xyz_lists), t.cat(d2_lists)
                ctx.features_merge(nF, k, nlists, xyz_all.data_ptr(), d2_all.data_ptr(), F.data_ptr(), self._p(self.props))
        finally:
            ix.free()
        self._mark("merge")
        self._mark("end")
        self.stats = {"halo_m": H, "own_in": int(n_own), "own": int(m), "halo_points": int(nl + nr), "boundary_queries": int(nF),
                      "sent": [0 if sd is None else int(sd[0].numel()) for sd in side],
                      "slab_columns": [P.bounds[self.rank], P.bounds[self.rank + 1]], "axis": "xy"[ax]}
        self.last_m = m
        return m

    def gather_gen(self, m):
        """The final result gather (north star): every rank's packed 24-byte records to rank 0, in rank order = slab
        order along x.  Returns (uint8 tensor, byte offsets) on rank 0, None elsewhere."""
        if self.packed is None:
            self.packed = self._alloc(self.cap * 24)
        self.ctx.props_pack(self._p(self.props), m, self._p(self.packed))
        if self.world == 1:
            return self.packed[:m * 24], np.array([0, m * 24], dtype=np.int64)
        return (yield ("gather", (self.packed, m * 24)))

    def gather_results(self, m):
        gen = self.gather_gen(m)
        return self.comm.run(gen) if self.world > 1 else self._drain(gen)

    @staticmethod
    def _drain(gen):
        try:
            next(gen)
        except StopIteration as e:
            return e.value
        raise RuntimeError("a single-rank step must not communicate")

    def close(self):
        self.src = self.cloud = self.part = self.work = self.props = self.packed = self.kth = None

    # ---- end to end: host buffers through the C ABI ----
    def time_e2e(self, steps, barrier, sync, keep=None, copy_only=False):
        """Input: this rank's range of the cloud, packed 32-byte records in pinned HOST memory; output: packed 24-byte
        LAS_POINT_PROPERTY records in pinned HOST memory (pipelined_e2e).  copy_only: the same uploads and downloads
        with no compute between them.  Returns (ms_per_step, h2d_bytes, d2h_bytes, featured per step)."""
        if self.packed is None:
            self.packed = self._alloc(self.cap * 24)
        packs = [self.packed, self._alloc(self.cap * 24)]
        m_fixed = self.last_m

        def run_step(dev_in, b):
            if copy_only:
                return self._p(packs[b]), m_fixed * 24, m_fixed
            m = self.step_device(dev_in)
            self.ctx.props_pack(self._p(self.props), m, self._p(packs[b]))
            return self._p(packs[b]), m * 24, m

        return pipelined_e2e(self.ctx, self._alloc, self._p(self.src), self.n * REC, self.cap * 24, run_step, steps, barrier,
                             sync, keep)


def pipelined_e2e(ctx, alloc, src_ptr, nb_in, out_cap, run_step, steps, barrier, sync, keep=None):
    """Times `steps` passes whose input starts in pinned HOST memory and whose results end in pinned HOST memory, every
    byte moved through the C ABI inside the timed region.  The passes are pipelined over the library's copy lanes
    (pcp_copy_async): while pass i is processed, the input of pass i+1 is uploaded and the results of pass i-1 are
    downloaded, each pass with its own device and host buffers.  Wall clock around all of it (fill and drain
    included), barrier + device synchronisation on both sides.
    run_step(dev_in_ptr, b) -> (device pointer of the results, bytes, units).
    Returns (ms_per_step, h2d_bytes, d2h_bytes, units per step); keep: list that receives a copy of the last
    pass's host results (tests)."""
    import time
    host_in, pin_in = ctx.pinned(nb_in)
    pins = [ctx.pinned(out_cap) for _ in range(2)]
    dev_in = [alloc(nb_in) for _ in range(2)]
    ctx.memcpy(pin_in, src_ptr, nb_in)
    UP, DOWN = 0, (1, 2)
    dp = lambda b: dev_in[b].data_ptr() if hasattr(dev_in[b], "data_ptr") else dev_in[b].ptr  # noqa: E731

    def run(nsteps):
        units = d2h = 0
        ctx.copy_async(UP, dp(0), pin_in, nb_in)
        for i in range(nsteps):
            b = i & 1
            ctx.copy_fence(UP)            # pass i reads dev_in[b]: wait for its upload
            ctx.copy_fence(DOWN[b])       # ... and writes result buffer b: wait for the download of pass i-2
            if i + 1 < nsteps:            # dev_in[1-b] was last read by pass i-1, already submitted
                ctx.copy_async(UP, dp(1 - b), pin_in, nb_in)
            out_ptr, nbytes, u = run_step(dp(b), b)
            ctx.copy_async(DOWN[b], pins[b][1], out_ptr, nbytes)
            units += u
            d2h = nbytes
        ctx.copy_wait(DOWN[0])
        ctx.copy_wait(DOWN[1])
        return units, d2h

    try:
        run(2)  # warms the arena, the lanes and both buffer sets
        sync()
        barrier()
        t0 = time.perf_counter()
        units, d2h = run(steps)
        sync()
        barrier()
        ms = (time.perf_counter() - t0) * 1e3 / steps
        if keep is not None:
            keep.append(pins[(steps - 1) & 1][0][:d2h].copy())
    finally:
        ctx.synchronize()
        ctx.free_pinned(pin_in)
        for _, p in pins:
            ctx.free_pinned(p)
    return ms, nb_in, d2h, units / steps


def digest(torch, work, props, m):
    """Order-independent digest of a result: the sum over the points of a mix of the voxelised coordinates and the
    record's normal / Distance / curvature bits (PointID, a local index, is left out), mod 2^64, and the count.
    A single-device run and an N-rank run of the same cloud must agree on it.  -> (digest as hex string, count)"""
    if m == 0:
        return "%016x" % 0, 0
    w = work[:m * REC].view(torch.int64).view(m, 4)
    p = props[:m * 48].view(torch.int64).view(m, 6)
    acc = 0
    for lo in range(0, m, 1 << 24):  # in pieces: the temporaries stay small beside a 200 M-point cloud
        ww, pp = w[lo:lo + (1 << 24)], p[lo:lo + (1 << 24)]
        h = ww[:, 0] * -7046029254386353131 + ww[:, 1] * -4658895280553007687 + ww[:, 2] * 7640891576956012809
        h = (h ^ (h >> 29)) * -4417276706812531889
        h = h + pp[:, 0] * 2685821657736338717 + pp[:, 1] * 1181783497276652981 + pp[:, 2] * -8663945395140668459 \
            + pp[:, 3] * 6364136223846793005
        h = (h ^ (h >> 32)) * -2401053089206453570
        acc = (acc + int(h.sum().item())) & 0xFFFFFFFFFFFFFFFF
    return "%016x" % acc, int(m)


def combine_digests(parts):
    """Digest of the union of disjoint results (the ranks' shards)."""
    return "%016x" % (sum(int(d, 16) for d, _ in parts) & 0xFFFFFFFFFFFFFFFF), sum(c for _, c in parts)
