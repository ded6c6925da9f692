"""CPU tests of the multi-rank plumbing: the neighbour exchange over torch.distributed (gloo, world_size 2 and 3)
and the lock-step driver used to emulate all ranks on one GPU."""
import os
import socket

import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q, host_group=False):
    import torch
    import torch.distributed as dist
    from pointcloudprocess_b200.sharded import DistComm
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # host_group: the message sizes travel over a second (gloo) group, as bench.py sets it up beside NCCL
    comm = DistComm(dist, torch, rank, world, torch.device("cpu"),
                    host_group=dist.new_group(backend="gloo") if host_group else None)
    ok = True
    for rnd in range(3):  # variable sizes, including empty messages
        mk = lambda dst: torch.full((rnd * 7 + rank + 1 if (rnd + dst) % 2 else 0,), 10 * rank + dst, dtype=torch.uint8)
        got_l, got_r = comm.exchange(mk(0) if rank > 0 else None, mk(1) if rank < world - 1 else None)
        if rank > 0:  # what rank-1 sent to its right neighbour (dst code 1)
            n = rnd * 7 + (rank - 1) + 1 if (rnd + 1) % 2 else 0
            ok &= got_l.numel() == n and bool((got_l == 10 * (rank - 1) + 1).all())
        else:
            ok &= got_l.numel() == 0
        if rank < world - 1:
            n = rnd * 7 + (rank + 1) + 1 if (rnd + 0) % 2 else 0
            ok &= got_r.numel() == n and bool((got_r == 10 * (rank + 1) + 0).all())
        else:
            ok &= got_r.numel() == 0
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, ok))


@pytest.mark.parametrize("world,host_group", [(2, False), (3, False), (2, True)])
def test_neighbour_exchange_gloo(world, host_group):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q, host_group)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(r, True) for r in range(world)]


def test_lockstep_driver_routes_messages():
    from pointcloudprocess_b200.sharded import run_lockstep

    def gen(rank, world):
        got = yield ("L%d" % rank if rank > 0 else None, "R%d" % rank if rank < world - 1 else None)
        assert got == ("R%d" % (rank - 1) if rank > 0 else None, "L%d" % (rank + 1) if rank < world - 1 else None)
        got = yield (None, None)
        assert got == (None, None)
        return rank * 10

    assert run_lockstep([gen(r, 4) for r in range(4)]) == [0, 10, 20, 30]


def test_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` (the CPU arm the driver times beside the GPU arm) needs no GPU and prints one JSON
    line with the contract's keys; its e2e object repeats the value with no host<->device traffic."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=600, cwd=root)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "points/s" and line["higher_is_better"] is True
    assert line["metric"] == "kNN(k=20)+feature points/sec" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] in ("port", "reference") and line["cpu_baseline"]["cores"] >= 1
    assert line["cpu_baseline"]["value"] == line["value"]
    assert line["e2e"]["value"] == line["value"] and line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
