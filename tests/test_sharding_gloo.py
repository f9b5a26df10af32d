"""CPU tests of the N>1 path: the (sequence, GOP) -> rank map and the single statistics
reduction, run with world_size 2 over gloo (bench.py uses the same
functions with the nccl backend)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from dcvc_b200 import sharding


def test_units_cover_job_once():
    units = sharding.work_units(8, 96, 32)
    assert len(units) == 24
    for world in (1, 2, 4, 8, 5):
        parts = sharding.assign_units(units, world)
        flat = [u for p in parts for u in p]
        assert sorted(flat) == sorted(units)
        loads = [sum(u.n_frames for u in p) for p in parts]
        assert max(loads) - min(loads) <= 32
    # config 3 of BASELINE.json: 24 units -> exactly 3 per GPU at 8 GPUs
    assert [len(p) for p in sharding.assign_units(units, 8)] == [3] * 8


def test_ragged_last_gop_and_errors():
    units = sharding.work_units(1, 70, 32)
    assert [(u.first_frame, u.n_frames) for u in units] == [(0, 32), (32, 32), (64, 6)]
    parts = sharding.assign_units(units, 2)
    assert sorted(sum(u.n_frames for u in p) for p in parts) == [32, 38]
    assert sharding.work_units(0, 10, 4) == []
    for bad in (lambda: sharding.work_units(1, 8, 0), lambda: sharding.assign_units([], 0)):
        try:
            bad()
        except ValueError:
            continue
        raise AssertionError("expected ValueError")


def _fake_code(unit):
    """Stand-in for coding one unit: deterministic (bits, sse, frames) from the unit id."""
    g = torch.Generator().manual_seed(1000 * unit.sequence + unit.gop)
    v = torch.rand(2, generator=g, dtype=torch.float64)
    return float(v[0]) * 1e6 * unit.n_frames, float(v[1]) * unit.n_frames, float(unit.n_frames)


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        units = sharding.work_units(3, 20, 8)
        mine = sharding.assign_units(units, world)[rank]
        acc = [0.0, 0.0, 0.0]
        for u in mine:
            acc = [a + b for a, b in zip(acc, _fake_code(u))]
        dist.barrier()
        sums, ms = sharding.reduce_stats(acc, 10.0 + rank)
        q.put((rank, sums, ms, len(mine)))
    finally:
        dist.destroy_process_group()


def test_two_ranks_reduce_to_single_process_totals():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    units = sharding.work_units(3, 20, 8)
    want = [0.0, 0.0, 0.0]
    for u in units:
        want = [a + b for a, b in zip(want, _fake_code(u))]
    assert sum(g[3] for g in got) == len(units)
    for _, sums, ms, _ in got:
        assert ms == 11.0                                   # max over ranks
        for a, b in zip(sums, want):
            assert abs(a - b) <= 1e-9 * max(1.0, abs(b))
