"""Sub-batch sweep of the pair matcher: pairs/s vs pairs per forward at 1024 / 2048 / 4096 keypoints.

    python tools/pairs_sweep.py [--profile]      # --profile: torch profiler kernel table for one shape
"""
import argparse
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from lgattn_pkg import pkg  # noqa: E402


def feats(b, n, dev, seed=0):
    g = torch.Generator(device=dev).manual_seed(seed)
    k = torch.rand(2, b, n, 2, device=dev, generator=g) * 2 - 1
    d = torch.randn(2, b, n, 256, device=dev, generator=g)
    d = (d / d.norm(dim=-1, keepdim=True)).half()
    return k[0], k[1], d[0], d[1]


def time_shape(matcher, b, n, total_pairs, dev):
    f = feats(b, n, dev)
    reps = max(1, total_pairs // b)
    for _ in range(2):
        matcher.match(*f)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        matcher.match(*f)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    fl = pkg.matcher.flops_per_pair(n, n)["total"]
    pps = reps * b / (ms * 1e-3)
    return {"kpts": n, "pairs_per_forward": b, "pairs_per_s": pps, "ms_per_pair": ms / (reps * b), "model_tflops": pps * fl / 1e12}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--profile", action="store_true")
    ap.add_argument("--kpts", type=int, nargs="*", default=[1024, 2048, 4096])
    ap.add_argument("--subs", type=int, nargs="*", default=[1, 2, 4, 8, 16, 32, 64])
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    matcher = pkg.PairMatcher(seed=0, device=dev)
    if args.profile:
        from torch.profiler import ProfilerActivity, profile
        eager = pkg.PairMatcher(seed=0, device=dev, use_graph=False)
        f = feats(8, 2048, dev)
        eager.match(*f)
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            eager.match(*f)
            torch.cuda.synchronize()
        print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=70))
        return
    for n in args.kpts:
        for b in args.subs:
            if b * n * n * 4 > 8e9:
                continue
            print(json.dumps(time_shape(matcher, b, n, 64, dev)), flush=True)
        matcher._plans.clear()
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
