"""GPU box: correctness spot check + TFLOP/s of the throughput kernel at the batched shapes
(use LG_ATTN_LIB=... to time a variant build)."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from lgattn_pkg import pkg
from oracle import attention_oracle as orc

errs = []
for (b, n, m) in [(80, 777, 1000), (80, 128, 128), (40, 300, 2047)]:
    q, k, v = orc.make_qkv(1, b, 4, n, m)
    ref = orc.attention_np(q[:2], k[:2], v[:2])
    tq, tk, tv = (torch.from_numpy(x).cuda() for x in (q, k, v))
    errs.append(float(orc.max_abs_err(pkg.attention(tq, tk, tv)[:2].float().cpu().numpy(), ref)))
    errs.append(float(orc.rel_err(pkg.attention(tq, tk, tv, out_dtype=torch.float32)[:2].cpu().numpy(), ref)))
print("errs", ["%.1e" % e for e in errs])
for (b, n) in [(64, 2048), (16, 4096), (128, 1024), (256, 512), (512, 256)]:
    q = torch.randn(b, 4, n, 64, device="cuda").half(); k = torch.randn_like(q); v = torch.randn_like(q); o = torch.empty_like(q)
    for _ in range(5): pkg.attention(q, k, v, out=o)
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): pkg.attention(q, k, v, out=o)
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 20)
    print(f"B{b} N{n}: {best*1e3:8.1f} us  {4.0*b*4*n*n*64/(best*1e-3)/1e12:7.1f} TFLOP/s")
