"""CPU tests of the multi-rank plumbing (pointcloudprocess_b200/sharded.py): every communication request of the step
over torch.distributed (gloo, world_size 2 and 3), the same requests through the lock-step driver that emulates all
ranks on one GPU, the slab cut, and the reference arm of bench.py."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _toy_step(torch, rank, world):
    """A generator that uses every request kind the real step uses and returns what it saw."""
    u8 = lambda n, v: torch.full((n,), v, dtype=torch.uint8)  # noqa: E731
    seen = {}
    g = yield ("allgather", torch.tensor([rank * 1.5, 7.0], dtype=torch.float64))
    seen["allgather"] = g.tolist()
    carry = torch.zeros(3, dtype=torch.float64)
    yield ("chain_recv", carry)
    carry += rank + 1                      # the running sum walks from rank to rank
    yield ("chain_send", carry)
    yield ("bcast", (carry, world - 1))
    seen["carry"] = carry.tolist()
    h = torch.arange(5, dtype=torch.int64) * (rank + 1)
    yield ("allreduce", h)
    seen["hist"] = h.tolist()
    # all-to-all: rank r sends (q + 1) bytes of value 10 r + q to rank q
    counts = [q + 1 for q in range(world)]
    buf = torch.cat([u8(q + 1, 10 * rank + q) for q in range(world)])
    out = torch.zeros(64, dtype=torch.uint8)
    nb = yield ("alltoall", (buf, counts, out))
    seen["a2a"] = (nb, out[:nb].tolist())
    # neighbour exchange with variable sizes, empty and missing messages
    left = u8(rank + 2, 100 + rank) if rank > 0 else None
    right = u8(0, 0) if rank % 2 else u8(3, 200 + rank)
    fl, fr = yield ("exchange", (left, right if rank < world - 1 else None))
    seen["exch"] = (None if fl is None else fl.tolist(), None if fr is None else fr.tolist())
    res = yield ("gather", (u8(8, rank + 1), rank + 2))
    seen["gather"] = None if res is None else (res[0].tolist(), [int(v) for v in res[1]])
    return seen


def _expected(rank, world):
    exp = {"allgather": [[q * 1.5, 7.0] for q in range(world)],
           "carry": [float(sum(q + 1 for q in range(world)))] * 3,
           "hist": [i * sum(q + 1 for q in range(world)) for i in range(5)]}
    a2a = []
    for q in range(world):
        a2a += [10 * q + rank] * (rank + 1)
    exp["a2a"] = (len(a2a), a2a)
    fl = None if rank == 0 else ([] if (rank - 1) % 2 else [200 + rank - 1] * 3)
    fr = None if rank == world - 1 else [100 + rank + 1] * (rank + 3)
    exp["exch"] = (fl, fr)
    if rank == 0:
        cat, off = [], [0]
        for q in range(world):
            cat += [q + 1] * (q + 2)
            off.append(len(cat))
        exp["gather"] = (cat, off)
    else:
        exp["gather"] = None
    return exp


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from pointcloudprocess_b200.sharded import DistComm
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    comm = DistComm(dist, torch, rank, world, torch.device("cpu"))
    ok = True
    for _ in range(2):  # twice: the persistent size words are reused
        got = comm.run(_toy_step(torch, rank, world))
        ok &= got == _expected(rank, world)
        if not ok:
            print("rank", rank, got, _expected(rank, world), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, ok))


@pytest.mark.parametrize("world", [2, 3])
def test_requests_over_gloo(world):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(r, True) for r in range(world)]


@pytest.mark.parametrize("world", [1, 2, 4])
def test_lockstep_driver_answers_the_same_requests(world):
    import torch
    from pointcloudprocess_b200.sharded import run_lockstep
    got = run_lockstep([_toy_step(torch, r, world) for r in range(world)])
    for r in range(world):
        assert got[r] == _expected(r, world), r


def test_slab_cut_keeps_every_slab_at_least_a_halo_wide():
    from pointcloudprocess_b200.sharded import cut_slabs
    cpb, leaf = 50, 0.02                      # 1 m bins
    hist = np.zeros(102, dtype=np.int64)
    hist[:100] = 1000
    b = cut_slabs(hist, 4, cpb, 100 * cpb - 1, leaf, 1.0)
    assert b[0] == 0 and b[-1] == 102 * cpb and all(b[i] < b[i + 1] for i in range(4))
    assert [x // cpb for x in b[1:-1]] == [25, 50, 75]   # equal counts
    # all the points in two bins: the cut would make metre-wide (or empty) slabs; with a 3 m halo every slab is >= 3 m
    hist = np.zeros(42, dtype=np.int64)
    hist[20:22] = 10 ** 6
    b = cut_slabs(hist, 4, cpb, 40 * cpb, leaf, 3.0)
    w = [(b[i + 1] - b[i]) * leaf for i in range(4)]
    assert min(w) >= 3.0 - 1e-9 and all(b[i] < b[i + 1] for i in range(4))
    with pytest.raises(ValueError):
        cut_slabs(np.ones(6, dtype=np.int64), 4, cpb, 4 * cpb, leaf, 3.0)  # 5 m of cloud cannot hold 4 slabs of 3 m


def test_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` (the CPU arm the driver times beside the GPU arm) needs no GPU, ignores
    OMP_NUM_THREADS (torchrun sets it to 1) and prints one JSON line with the contract's keys."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, OMP_NUM_THREADS="1")
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                          "--sample", "200000"], capture_output=True, text=True, timeout=600, cwd=root, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "points/s" and line["higher_is_better"] is True
    assert line["metric"] == "kNN(k=20)+feature points/sec" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] in ("port", "reference")
    assert line["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))
    assert line["cpu_baseline"]["value"] == line["value"]
    assert line["e2e"]["value"] == line["value"] and line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    for cfg in ("1", "2"):  # the other configs' CPU arms (numpy generators of the strip and the forest)
        out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup",
                              "0", "--config", cfg, "--sample", "100000"], capture_output=True, text=True, timeout=600, cwd=root)
        assert out.returncode == 0, out.stderr[-2000:]
        assert json.loads(out.stdout.strip().splitlines()[-1])["value"] > 0
