"""GPU box: run one full-grid kernel (lg_attn_debug_set_throughput_kernel mode) on one shape — the command ncu wraps.

    python tools/run_mode.py MODE B N [ITERS]
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from lgattn_pkg import pkg  # noqa: E402

mode, b, n = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 3
lib = pkg.load()
if mode >= 0:
    assert lib.lg_attn_debug_set_throughput_kernel(mode) == 0
q = torch.randn(b, 4, n, 64, device="cuda").half()
k = torch.randn_like(q)
v = torch.randn_like(q)
o = torch.empty_like(q)
for _ in range(iters):
    pkg.attention(q, k, v, out=o)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters):
    pkg.attention(q, k, v, out=o)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
print(f"mode {mode} B{b} N{n}: {ms * 1e3:.1f} us  {4.0 * b * 4 * n * n * 64 / (ms * 1e-3) / 1e12:.1f} TFLOP/s")
