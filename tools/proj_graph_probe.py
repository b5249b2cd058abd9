"""GPU box: the input projection inside a CUDA graph, the way the matcher runs it — a chain of dependent launches on ONE output
buffer (so neighbours cannot overlap): library GEMM + lg_match_rotary against lg_match_proj_rotary, us per projection.

    python tools/proj_graph_probe.py [rows ...]
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from lgattn_pkg import pkg  # noqa: E402
import tools.perf_compare as pc  # noqa: E402

M = pkg.matcher
dev = "cuda:0"
for rows in [int(a) for a in sys.argv[1:]] or [2048, 4096, 8192, 16384]:
    for n_blocks, n_rot in ((3, 2), (2, 0)):
        n = 256 * n_blocks
        xc = torch.randn(rows, 512, device=dev).half()
        x = xc[:, :256]
        w = (torch.randn(n, 256, device=dev) / 16).half()
        w_t = w.t().contiguous()
        bias = torch.randn(n, device=dev).half()
        cs = M.posenc_table(torch.rand(rows, 2, device=dev) * 2 - 1, torch.randn(32, 2, device=dev))
        out = torch.empty(rows, n, device=dev, dtype=torch.float16)

        def old(i, s):
            torch.addmm(bias, x, w_t, out=out)
            if n_rot:
                M.rotary_(out, n_rot, cs)

        def new(i, s):
            M.proj_rotary(x, w, bias, n_rot, cs if n_rot else None, out)

        t_old, t_new = pc.time_graph_chain(old, 32), pc.time_graph_chain(new, 32)
        print(f"rows {rows:6d} x {n}: GEMM{' + rotary' if n_rot else ''} {t_old:6.2f} us | fused {t_new:6.2f} us", flush=True)
