"""Debug aid (GPU box): run the kernel on a few shapes and print error statistics vs the oracle.

    python tools/gpu_check.py [--quick]
"""
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from lgattn_pkg import pkg  # noqa: E402
from oracle import attention_oracle as orc  # noqa: E402


def ref_gpu(q, k, v):
    q, k, v = q.double(), k.double(), v.double()
    s = (q / 8) @ k.transpose(-1, -2)
    return (s.softmax(-1) @ v)


def run(b, h, n, m, mode, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    in_dt = torch.float32 if mode == "f32_f32" else torch.float16
    out_dt = torch.float16 if mode == "f16_f16" else torch.float32
    q = torch.randn(b, h, n, 64, device="cuda", generator=g).to(in_dt)
    k = torch.randn(b, h, m, 64, device="cuda", generator=g).to(in_dt)
    v = torch.randn(b, h, m, 64, device="cuda", generator=g).to(in_dt)
    o = pkg.attention(q, k, v, out_dtype=out_dt)
    torch.cuda.synchronize()
    if mode == "f32_f32":
        r = ref_gpu(q.half(), k.half(), v.half())
    else:
        r = ref_gpu(q, k, v)
    err = (o.double() - r).abs()
    bad = torch.isnan(o).sum().item()
    mx = err.max().item()
    rel = mx / r.abs().max().item()
    print(f"{mode} B{b} H{h} N{n} M{m}: max_abs={mx:.3e} rel={rel:.3e} nan={bad} "
          f"mean_abs={err.mean().item():.3e}", flush=True)
    if mx > 5e-3 or bad:
        # where is it wrong? per (row-block of 32, col-block of 8) summary for head 0
        e = err[0, 0]
        rows = e.max(dim=1).values
        worst = torch.topk(rows, min(5, rows.numel()))
        print("   worst rows:", worst.indices.tolist(), [f"{x:.2e}" for x in worst.values.tolist()])
        cols = e.max(dim=0).values
        print("   col err  :", [f"{x:.1e}" for x in cols[:16].tolist()])
        print("   got[0,0,0,:8]", o[0, 0, 0, :8].tolist())
        print("   ref[0,0,0,:8]", r[0, 0, 0, :8].tolist())
    return mx


def main():
    print(torch.cuda.get_device_name(0), torch.version.cuda, flush=True)
    shapes = [(1, 1, 128, 128), (1, 4, 128, 256), (1, 4, 1024, 1024), (1, 4, 777, 1000), (1, 4, 1, 1),
              (2, 4, 65, 129), (1, 4, 513, 63)]
    if "--quick" in sys.argv:
        shapes = shapes[:3]
    for mode in ("f16_f16", "f16_f32", "f32_f32"):
        for (b, h, n, m) in shapes:
            run(b, h, n, m, mode)
    # crude timing
    for (b, n) in ((1, 1024), (64, 2048)):
        q = torch.randn(b, 4, n, 64, device="cuda").half()
        k = torch.randn(b, 4, n, 64, device="cuda").half()
        v = torch.randn(b, 4, n, 64, device="cuda").half()
        o = torch.empty_like(q)
        for _ in range(5):
            pkg.attention(q, k, v, out=o)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 50
        e0.record()
        for _ in range(iters):
            pkg.attention(q, k, v, out=o)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        print(f"timing B{b} N=M={n}: {ms*1e3:.1f} us  {orc.flops(b,4,n,n)/ms/1e9:.1f} TFLOP/s", flush=True)


if __name__ == "__main__":
    main()
