"""CPU tests of the N>1 host logic: pair sharding, max-over-ranks timing and the host gather, with
two real processes over the gloo backend (the GPU box uses the same code over NCCL)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from lgattn_pkg import pkg


def test_round_robin_partition_is_exact():
    sh = pkg.sharding
    for total, world in [(0, 1), (1, 8), (64, 8), (65, 4), (512, 8), (7, 3)]:
        owned = [sh.pairs_of_rank(total, world, r) for r in range(world)]
        flat = sorted(p for o in owned for p in o)
        assert flat == list(range(total))
        assert max(len(o) for o in owned) - min(len(o) for o in owned) <= 1
        assert all(p % world == r for r, o in enumerate(owned) for p in o)
    with pytest.raises(ValueError):
        sh.pairs_of_rank(4, 2, 2)


def test_sub_batches_cover_local_pairs():
    sh = pkg.sharding
    assert sh.sub_batches(64, 32) == [(0, 32), (32, 64)]
    assert sh.sub_batches(10, 4) == [(0, 4), (4, 8), (8, 10)]
    assert sh.sub_batches(0, 8) == []


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, total: int, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sh = pkg.sharding
    mine = sh.pairs_of_rank(total, world, rank)
    # stand-in for the GPU matcher: a deterministic per-pair "match count"
    local = {p: (p * 7919) % 1024 for p in mine}
    t = sh.max_over_ranks(dist, 0.5 + rank)          # the slowest rank defines the step time
    merged = sh.gather_to_host(dist, local)
    dist.barrier()
    q.put((rank, t, None if merged is None else sorted(merged.items())))
    dist.destroy_process_group()


def test_two_ranks_gloo_shard_time_and_gather():
    world, total = 2, 13
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [g[1] for g in got] == [1.5, 1.5]
    assert got[1][2] is None
    assert got[0][2] == [(p, (p * 7919) % 1024) for p in range(total)]


def test_single_process_paths_need_no_process_group():
    sh = pkg.sharding
    assert sh.max_over_ranks(dist, 0.25) == 0.25
    assert sh.gather_to_host(dist, {0: 1}) == {0: 1}


def test_cpulist_parsing_and_gpu_local_binding(tmp_path):
    import os

    from lgattn_pkg import pkg

    sh = pkg.sharding
    assert sh.parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert sh.parse_cpulist("") == set()
    # no sysfs entry: reported, not raised
    assert sh.bind_host_to_gpu("0000:ff:00.0", sysfs_root=str(tmp_path))["bound"] is False
    # a fake device whose local CPUs are exactly one CPU this process may use
    before = os.sched_getaffinity(0)
    cpu = min(before)
    d = tmp_path / "0000:1b:00.0"
    d.mkdir()
    (d / "local_cpulist").write_text(f"{cpu}\n")
    (d / "numa_node").write_text("1\n")
    try:
        r = sh.bind_host_to_gpu("1B:00.0", sysfs_root=str(tmp_path))
        assert r["bound"] is True and r["numa_node"] == 1
        assert os.sched_getaffinity(0) == {cpu}
    finally:
        os.sched_setaffinity(0, before)
    # local CPUs outside the cpuset: left alone
    (d / "local_cpulist").write_text("100000\n")
    assert sh.bind_host_to_gpu("0000:1b:00.0", sysfs_root=str(tmp_path))["bound"] is False
    assert os.sched_getaffinity(0) == before
