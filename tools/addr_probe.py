"""GPU box: is the B128 N=M=1024 throughput sensitive to where the tensors sit? (bimodal 670 / 750 TFLOP/s across runs)"""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from lgattn_pkg import pkg
b, n = 128, 1024
nel = b * 4 * n * 64
def run(q, k, v, o, iters=20):
    for _ in range(3): pkg.attention(q, k, v, out=o)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters): pkg.attention(q, k, v, out=o)
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / iters)
    return 4.0 * b * 4 * n * n * 64 / (best * 1e-3) / 1e12
pool = torch.empty(4 * nel + (64 << 20), dtype=torch.float16, device="cuda")
pool.normal_()
for mode in (-1, 0, 1):
    pkg.load().lg_attn_debug_set_throughput_kernel(mode)
    res = []
    for off_kb in (0, 4, 64, 512, 1024, 2048, 4096, 9000):
        off = off_kb * 512  # halves
        q = pool[off:off + nel].view(b, 4, n, 64); k = pool[off + nel:off + 2 * nel].view(b, 4, n, 64)
        v = pool[off + 2 * nel:off + 3 * nel].view(b, 4, n, 64); o = pool[off + 3 * nel:off + 4 * nel].view(b, 4, n, 64)
        res.append(round(run(q, k, v, o)))
    print("mode", mode, "offsets KiB (0,4,64,512,1024,2048,4096,9000):", res)
    # separate allocations
    res = []
    for _ in range(4):
        q = torch.randn(b, 4, n, 64, device="cuda").half(); k = torch.randn_like(q); v = torch.randn_like(q); o = torch.empty_like(q)
        res.append(round(run(q, k, v, o)))
        junk = torch.empty(37 << 20, dtype=torch.uint8, device="cuda")
    print("   fresh allocations:", res)
