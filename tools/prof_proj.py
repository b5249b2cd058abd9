"""Profiling driver (run under ncu): three launches of lg_match_proj_rotary at 262144 rows x 768 columns, rotary on q | k."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from lgattn_pkg import pkg  # noqa: E402

M = pkg.matcher
dev = "cuda:0"
rows = 262144
x = torch.randn(rows, 512, device=dev).half()[:, :256]
w = (torch.randn(768, 256, device=dev) / 16).half()
bias = torch.randn(768, device=dev).half()
cs = M.posenc_table(torch.rand(rows, 2, device=dev) * 2 - 1, torch.randn(32, 2, device=dev))
out = torch.empty(rows, 768, device=dev, dtype=torch.float16)
for _ in range(3):
    M.proj_rotary(x, w, bias, 2, cs, out)
torch.cuda.synchronize()
print("ok")
