"""Pin the ceiling of the op-level e2e leg with a measurement: N ranks (one per GPU) each loop the bench's per-problem copies —
H2D 1.5 MiB (q|k|v of one B1 H4 N=M=1024 FP16 problem) + D2H 0.5 MiB (o) — from the same kind of pinned buffers on 4 streams, with
NO kernels.  If the aggregate bytes/s here equals what bench.py's e2e leg moves at the same N, the e2e number is the box's
host-memory / PCIe ceiling and not the session's submit path.

    python tools/copy_only_probe.py --gpus 8        # spawns itself once per rank count 1, 2, 4, 8 (<= --gpus)
"""
import argparse
import json
import os
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
H2D, D2H, CALLS, FLOPS = 3 * 4 * 1024 * 64 * 2, 4 * 1024 * 64 * 2, 128, 4.0 * 4 * 1024 * 1024 * 64


def worker(rank: int, world: int, steps: int, group: int = 1) -> None:
    import torch

    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    src = torch.empty(CALLS, H2D, dtype=torch.uint8).pin_memory()
    dst_h = torch.empty(CALLS, D2H, dtype=torch.uint8).pin_memory()
    slots = 4
    d_in = [torch.empty(group, H2D, dtype=torch.uint8, device=dev) for _ in range(slots)]
    d_out = [torch.empty(group, D2H, dtype=torch.uint8, device=dev) for _ in range(slots)]
    streams = [torch.cuda.Stream(dev) for _ in range(slots)]

    def step():   # `group` problems per transfer each way (bench.py's e2e leg: LG_E2E_GROUP)
        for j, i in enumerate(range(0, CALLS, group)):
            s = j % slots
            with torch.cuda.stream(streams[s]):
                d_in[s].copy_(src[i:i + group], non_blocking=True)
                dst_h[i:i + group].copy_(d_out[s], non_blocking=True)
        torch.cuda.synchronize(dev)

    for _ in range(3):
        step()
    # crude cross-process start barrier: a file per rank
    flag = Path(os.environ["PROBE_DIR"]) / f"ready_{world}_{rank}"
    flag.write_text("1")
    while len(list(Path(os.environ["PROBE_DIR"]).glob(f"ready_{world}_*"))) < world:
        time.sleep(0.001)
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = time.perf_counter() - t0
    print(json.dumps({"rank": rank, "seconds": dt, "problems": CALLS * steps}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--worker", type=int, default=-1)
    ap.add_argument("--world", type=int, default=1)
    ap.add_argument("--group", type=int, default=4, help="problems per transfer (bench.py moves LG_E2E_GROUP = 4 per copy)")
    args = ap.parse_args()
    if args.worker >= 0:
        worker(args.worker, args.world, args.steps, args.group)
        return
    import tempfile

    for world in [n for n in (1, 2, 4, 8) if n <= args.gpus]:
        with tempfile.TemporaryDirectory() as d:
            env = dict(os.environ, PROBE_DIR=d)
            procs = [subprocess.Popen([sys.executable, __file__, "--worker", str(r), "--world", str(world), "--steps", str(args.steps), "--group", str(args.group)],
                                      env=env, stdout=subprocess.PIPE, text=True) for r in range(world)]
            outs = [json.loads(p.communicate()[0].strip().splitlines()[-1]) for p in procs]
        slowest = max(o["seconds"] for o in outs)
        problems = sum(o["problems"] for o in outs)
        gbs = problems * (H2D + D2H) / slowest / 1e9
        print(json.dumps({"n_gpus": world, "problems_per_transfer": args.group, "copy_only_GBps_aggregate": round(gbs, 1), "h2d_GBps": round(problems * H2D / slowest / 1e9, 1),
                          "d2h_GBps": round(problems * D2H / slowest / 1e9, 1), "problems_per_s": round(problems / slowest),
                          "equivalent_e2e_tflops_if_copy_bound": round(problems / slowest * FLOPS / 1e12, 1)}), flush=True)


if __name__ == "__main__":
    main()
