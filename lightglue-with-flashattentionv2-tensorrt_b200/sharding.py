"""Host-side sharding of independent image pairs over the GPUs of one box.

Pairs are independent (``demo/demo_mono.cpp:211-412`` processes them one by one), so the only
multi-GPU logic is bookkeeping: pair ``p`` goes to rank ``p mod G``; every rank matches its own pairs
on its own GPU and stream; per-pair results (match counts / indices) are gathered to rank 0 on the
HOST.  There is no collective in the data path — ``torch.distributed`` (NCCL on the GPU box, gloo in
the CPU tests) is used for the barrier, the max-over-ranks step time and the host gather only.
"""
from __future__ import annotations


def pairs_of_rank(total_pairs: int, world: int, rank: int) -> list[int]:
    """Global pair indices owned by ``rank`` (round-robin: pair p -> rank p mod world)."""
    if world <= 0 or not (0 <= rank < world) or total_pairs < 0:
        raise ValueError("bad world / rank / total_pairs")
    return list(range(rank, total_pairs, world))


def sub_batches(n_pairs: int, pairs_per_forward: int) -> list[tuple[int, int]]:
    """[start, stop) slices of a rank's local pairs, ``pairs_per_forward`` at a time (last may be short)."""
    if pairs_per_forward <= 0:
        raise ValueError("pairs_per_forward must be positive")
    return [(s, min(s + pairs_per_forward, n_pairs)) for s in range(0, n_pairs, pairs_per_forward)]


def max_over_ranks(dist, seconds: float, device=None) -> float:
    """Step time of the job = the slowest rank's (all_reduce MAX); identity when not initialised."""
    import torch

    if dist is None or not dist.is_available() or not dist.is_initialized():
        return float(seconds)
    t = torch.tensor([seconds], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_to_host(dist, local: dict) -> dict | None:
    """Gather ``{global_pair_index: result}`` dicts to rank 0 (host objects); returns the merged dict on
    rank 0, None elsewhere.  Raises if two ranks claim the same pair."""
    if dist is None or not dist.is_available() or not dist.is_initialized():
        return dict(local)
    world, rank = dist.get_world_size(), dist.get_rank()
    bucket = [None] * world if rank == 0 else None
    dist.gather_object(local, bucket, dst=0)
    if rank != 0:
        return None
    merged: dict = {}
    for part in bucket:
        for k, v in part.items():
            if k in merged:
                raise RuntimeError(f"pair {k} was matched by two ranks")
            merged[k] = v
    return merged


def parse_cpulist(text: str) -> set[int]:
    """``"0-3,8,10-11"`` (the sysfs ``local_cpulist`` format) -> {0, 1, 2, 3, 8, 10, 11}."""
    cpus: set[int] = set()
    for part in text.strip().split(","):
        part = part.strip()
        if not part:
            continue
        lo, _, hi = part.partition("-")
        a, b = int(lo), int(hi or lo)
        if b < a:
            raise ValueError(f"bad cpu range {part!r}")
        cpus.update(range(a, b + 1))
    return cpus


def bind_host_to_gpu(pci_bus_id: str, sysfs_root: str = "/sys/bus/pci/devices") -> dict:
    """Pin the calling process to the CPUs local to one GPU (``<sysfs>/<pci>/local_cpulist``).

    One process drives one GPU; its pinned staging buffers are first-touched after this call and so
    land on the GPU's own NUMA node — host<->device copies of 8 ranks then stay off the socket
    interconnect.  Best effort: returns ``{"bound": False, "why": ...}`` when sysfs has no entry, the
    list does not intersect the cpuset this process may use, or the kernel refuses.
    """
    import os

    dev = pci_bus_id.lower()
    if dev.count(":") == 1:
        dev = "0000:" + dev
    path = os.path.join(sysfs_root, dev, "local_cpulist")
    try:
        with open(path) as f:
            local = parse_cpulist(f.read())
    except (OSError, ValueError) as exc:
        return {"bound": False, "why": f"{path}: {exc.__class__.__name__}"}
    try:
        allowed = os.sched_getaffinity(0)
    except (AttributeError, OSError):
        return {"bound": False, "why": "sched_getaffinity unavailable"}
    target = local & allowed
    if not target:
        return {"bound": False, "why": "local cpus not in this process's cpuset"}
    if target == allowed:
        return {"bound": True, "cpus": len(target), "numa_node": _read_int(os.path.join(sysfs_root, dev, "numa_node")),
                "note": "already local"}
    try:
        os.sched_setaffinity(0, target)
    except OSError as exc:
        return {"bound": False, "why": f"sched_setaffinity: {exc.__class__.__name__}"}
    return {"bound": True, "cpus": len(target), "numa_node": _read_int(os.path.join(sysfs_root, dev, "numa_node"))}


def _read_int(path: str):
    try:
        with open(path) as f:
            return int(f.read().strip())
    except (OSError, ValueError):
        return None
