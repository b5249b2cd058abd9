"""GPU box: the fused projection + rotary kernel (lg_match_proj_rotary) against the path it replaces (library GEMM + lg_match_rotary),
same tensors, CUDA events, best of `reps` bursts of `iters` launches.

    python tools/proj_probe.py [rows ...]
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
    for n_blocks, n_rot in ((3, 2), (3, 0), (2, 0)):
        n = 256 * n_blocks
        x = torch.randn(rows, 512, device=dev).half()[:, :256]
        w = (torch.randn(n, 256, device=dev) / 16).half()
        w_t = w.t().contiguous()
        bias = torch.randn(n, device=dev).half()
        kpts = torch.rand(rows, 2, device=dev) * 2 - 1
        wr = torch.randn(32, 2, device=dev)
        cs = M.posenc_table(kpts, wr)
        o0 = torch.empty(rows, n, device=dev, dtype=torch.float16)
        o1 = torch.empty_like(o0)

        def old():
            torch.addmm(bias, x, w_t, out=o0)
            if n_rot:
                M.rotary_(o0, n_rot, cs)

        def new():
            M.proj_rotary(x, w, bias, n_rot, cs if n_rot else None, o1)

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
        diff = float((o0.float() - o1.float()).abs().max())
        t_old, t_new = timeit(old), timeit(new)
        gb = (rows * (512 + (256 if n_rot else 0)) + rows * n * 2) / 1e9
        print(f"rows {rows:7d} x {n:3d} cols (rotary blocks {n_rot}): GEMM{' + rotary' if n_rot else ''} {t_old:8.1f} us | fused {t_new:8.1f} us "
              f"= {gb / t_new * 1e6:6.0f} GB/s algorithmic | speed-up {t_old / t_new:5.2f}x | max |fused - unfused| {diff:.4f}", flush=True)
