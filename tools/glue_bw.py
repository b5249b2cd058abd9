"""GPU box: achieved bandwidth of the HBM-bound glue kernels at the pair-matcher's bench sizes
(64 pairs x 1024 keypoints and 32 pairs x 2048: 131072 rows per launch)."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from lgattn_pkg import pkg
from importlib import import_module
M = pkg.matcher

def timeit(fn, iters=50):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters

for rows in (16384, 32768, 131072):
    h = torch.randn(rows, 512, device="cuda").half()
    g = torch.ones(512, device="cuda"); b = torch.zeros(512, device="cuda")
    us = timeit(lambda: M.layernorm_gelu_(h, g, b))
    print(f"layernorm_gelu rows={rows}: {us:7.1f} us  {rows*2048/us/1e3:7.1f} GB/s (read+write)")
    proj = torch.randn(rows, 768, device="cuda").half()
    cs = torch.randn(rows, 32, 2, device="cuda")
    us = timeit(lambda: M.rotary_(proj, 2, cs))
    print(f"rotary         rows={rows}: {us:7.1f} us  {rows*(2*512*2+256)/us/1e3:7.1f} GB/s (read+write q,k + cos/sin)")
