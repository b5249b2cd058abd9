"""GPU box: launch time of the 4-Q-tile throughput kernel against the number of waves of its persistent grid (one item = 4 Q tiles of
one (batch, head)) — slope = time per item, intercept = fixed cost per launch (prologue, pipeline fill, drain).

    python tools/wave_probe.py [N] [mode]
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from lgattn_pkg import pkg  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
mode = int(sys.argv[2]) if len(sys.argv) > 2 else 1
lib = pkg.load()
assert lib.lg_attn_debug_set_throughput_kernel(mode) == 0
groups = (n + 511) // 512   # items per (batch, head) for the 4-tile layout
for waves in (1, 2, 3, 4, 6, 8, 12):
    bh = 148 * waves // groups
    b = bh // 4
    q = torch.randn(b, 4, n, 64, device="cuda").half()
    k = torch.randn_like(q)
    v = torch.randn_like(q)
    o = torch.empty_like(q)
    for _ in range(5):
        pkg.attention(q, k, v, out=o)
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            pkg.attention(q, k, v, out=o)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 20)
    items = b * 4 * groups
    tf = 4.0 * b * 4 * n * n * 64 / (best * 1e-3) / 1e12
    print(f"N={n} B={b} items={items} waves={items / 148:.2f}  {best * 1e3:8.1f} us/launch  {tf:7.1f} TFLOP/s  [{pkg._binding.describe_launch(b, 4, n, n)}]", flush=True)
