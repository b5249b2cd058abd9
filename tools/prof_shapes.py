"""Profiling driver (run under ncu on the GPU box): a few launches of the attention kernel at the
latency shape (B1 N=M=1024) and the batched shape (B64 N=M=2048), FP16 out then FP32 out."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from lgattn_pkg import pkg  # noqa: E402

for (b, n) in ((1, 1024), (64, 2048)):
    q = torch.randn(b, 4, n, 64, device="cuda").half()
    k = torch.randn(b, 4, n, 64, device="cuda").half()
    v = torch.randn(b, 4, n, 64, device="cuda").half()
    for out_dt in (torch.float16, torch.float16, torch.float32):
        pkg.attention(q, k, v, out_dtype=out_dt)
    torch.cuda.synchronize()
print("ok")
# third shape: full grid with few KV tiles per Q tile -> persistent throughput kernel w4p
q = torch.randn(128, 4, 1024, 64, device="cuda").half(); k = torch.randn_like(q); v = torch.randn_like(q)
pkg.attention(q, k, v)
torch.cuda.synchronize()
print("ok w4p")
