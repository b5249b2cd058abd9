"""GPU box: the fused ffn-up kernel (lg_match_ffn1_ln_gelu) against the path it replaces (library GEMM + lg_match_layernorm_gelu),
same tensors, CUDA events, best of `reps` bursts of `iters` launches.

    python tools/ffn_probe.py [rows ...]
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from lgattn_pkg import pkg  # noqa: E402

M = pkg.matcher
rows_list = [int(a) for a in sys.argv[1:]] or [2048, 32768, 131072, 262144]
dev = "cuda:0"
for rows in rows_list:
    x = torch.randn(rows, 512, device=dev).half()
    w = (torch.randn(512, 512, device=dev) / 22.6).half()
    w_t = w.t().contiguous()
    bias = torch.randn(512, device=dev).half()
    gamma = torch.randn(512, device=dev)
    beta = torch.randn(512, device=dev)
    h0 = torch.empty(rows, 512, device=dev, dtype=torch.float16)
    h1 = torch.empty_like(h0)

    def old():
        torch.addmm(bias, x, w_t, out=h0)
        M.layernorm_gelu_(h0, gamma, beta)

    def gemm_only():
        torch.addmm(bias, x, w_t, out=h0)

    def new():
        M.ffn1_ln_gelu(x, w, bias, gamma, beta, h1)

    def timeit(f, iters=20, reps=5):
        for _ in range(3):
            f()
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                f()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / iters * 1e3)
        return best

    old()
    new()
    torch.cuda.synchronize()
    diff = float((h0.float() - h1.float()).abs().max())
    t_old, t_gemm, t_new = timeit(old), timeit(gemm_only), timeit(new)
    fl = 2.0 * rows * 512 * 512
    print(f"rows {rows:7d}: GEMM + LN/GELU {t_old:8.1f} us (GEMM alone {t_gemm:7.1f}) | fused {t_new:8.1f} us = {fl / t_new * 1e-6:6.0f} TFLOP/s "
          f"| speed-up {t_old / t_new:5.2f}x | max |fused - unfused| {diff:.4f}", flush=True)
