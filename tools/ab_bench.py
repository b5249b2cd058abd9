"""GPU box: interleaved A/B timing of library variants x full-grid kernel modes (+ torch SDPA) on the same tensors.
Variants alternate inside every repetition so that clock / power drift hits all of them alike; prints median and best TFLOP/s.

    python tools/ab_bench.py [--shapes 64x2048,128x1024] [--reps 7] [--iters 10] tag:mode[,mode...] ...   (tag 'main' = product build)
"""
import argparse
import ctypes
import json
import statistics
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "lightglue-with-flashattentionv2-tensorrt_b200" / "csrc"


def load(tag):
    path = CSRC / "build" / ("liblightglue_attention_plugin.so" if tag == "main" else f"variants/{tag}/liblightglue_attention_plugin.so")
    lib = ctypes.CDLL(str(path), mode=ctypes.RTLD_LOCAL)
    lib.lg_attn_hd64_launch.restype = ctypes.c_int
    lib.lg_attn_hd64_launch.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_int] * 5 + [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
    lib.lg_attn_debug_set_throughput_kernel.argtypes = [ctypes.c_int]
    assert lib.lg_attn_hd64_init_device(0) == 0
    return lib


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="64x2048,64x4096,128x1024,16x4096")
    ap.add_argument("--reps", type=int, default=7)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--out", default="gpurun_out/ab_bench.jsonl")
    ap.add_argument("specs", nargs="+")
    args = ap.parse_args()
    cands = [("sdpa", None, None)]
    for spec in args.specs:
        tag, _, modes = spec.partition(":")
        lib = load(tag)
        for m in (modes or "-1").split(","):
            cands.append((f"{tag}:{m}", lib, int(m)))
    stream = torch.cuda.current_stream().cuda_stream
    for shp in args.shapes.split(","):
        b, n = (int(x) for x in shp.split("x"))
        q = torch.randn(b, 4, n, 64, device="cuda").half()
        k = torch.randn_like(q)
        v = torch.randn_like(q)
        o = torch.empty_like(q)
        ref = torch.nn.functional.scaled_dot_product_attention(q[:1], k[:1], v[:1]).float()
        flops = 4.0 * b * 4 * n * n * 64

        def run(c):
            name, lib, mode = c
            if lib is None:
                torch.nn.functional.scaled_dot_product_attention(q, k, v)
            else:
                lib.lg_attn_debug_set_throughput_kernel(mode)
                rc = lib.lg_attn_hd64_launch(q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), b, 4, n, n, 0, None, 0, stream)
                assert rc == 0, rc

        times = {c[0]: [] for c in cands}
        errs = {}
        for c in cands:
            for _ in range(3):
                run(c)
            torch.cuda.synchronize()
            if c[1] is not None:
                errs[c[0]] = float((o[:1].float() - ref).abs().max())
        for rep in range(args.reps):
            for c in cands:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(args.iters):
                    run(c)
                e1.record()
                torch.cuda.synchronize()
                times[c[0]].append(e0.elapsed_time(e1) / args.iters)
        row = {"shape": f"B{b}_N{n}"}
        for name, ts in times.items():
            row[name] = {"med": round(flops / (statistics.median(ts) * 1e-3) / 1e12, 1), "best": round(flops / (min(ts) * 1e-3) / 1e12, 1)}
            if name in errs:
                row[name]["err"] = float(f"{errs[name]:.1e}")
        print(json.dumps(row), flush=True)
        with open(ROOT / args.out, "a") as f:
            f.write(json.dumps(row) + "\n")


if __name__ == "__main__":
    main()
