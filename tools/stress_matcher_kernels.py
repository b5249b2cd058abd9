"""GPU box: randomised stress of the matcher's tcgen05 kernels — random row counts, both ffn variants (must agree bit for bit with
each other, and with themselves when repeated), projection against an FP32 torch restatement; repeated launches back to back to
shake out ordering bugs (barrier phases across tiles, staging reuse, PDL).

    python tools/stress_matcher_kernels.py [iterations]
"""
import random
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from lgattn_pkg import pkg  # noqa: E402

M, lib, dev = pkg.matcher, pkg.load(), "cuda:0"
rng = random.Random(1234)
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 40
w1 = (torch.randn(512, 512, device=dev) / 22.6).half()
b1 = torch.randn(512, device=dev).half()
gamma, beta = torch.randn(512, device=dev), torch.randn(512, device=dev)
wq = (torch.randn(768, 256, device=dev) / 16).half()
bq = torch.randn(768, device=dev).half()
wr = torch.randn(32, 2, device=dev)
worst = 0.0
for it in range(iters):
    rows = rng.choice([rng.randint(1, 600), rng.randint(600, 20000), rng.randint(20000, 90000)])
    x = torch.randn(rows, 512, device=dev).half()
    outs = []
    for pair in (0, 1, 0, 1):
        lib.lg_match_debug_set_ffn_pair(pair)
        o = torch.empty(rows, 512, device=dev, dtype=torch.float16)
        for _ in range(3):   # back to back on the same output
            M.ffn1_ln_gelu(x, w1, b1, gamma, beta, o)
        outs.append(o)
    lib.lg_match_debug_set_ffn_pair(-1)
    torch.cuda.synchronize()
    assert all(torch.equal(outs[0], t) for t in outs[1:]), f"ffn variants disagree at rows={rows}"
    ref = torch.nn.functional.gelu(torch.nn.functional.layer_norm((x.float() @ w1.float().t() + b1.float()).half().float(), (512,), gamma, beta, 1e-5))
    e = float((outs[0].float() - ref).abs().max())
    assert e <= 2e-2, (rows, e)
    cs = M.posenc_table(torch.rand(rows, 2, device=dev) * 2 - 1, wr)
    p1, p2 = torch.empty(rows, 768, device=dev, dtype=torch.float16), torch.empty(rows, 768, device=dev, dtype=torch.float16)
    for _ in range(3):
        M.proj_rotary(x[:, :256], wq, bq, 2, cs, p1)
    torch.addmm(bq, x[:, :256], wq.t().contiguous(), out=p2)
    M.rotary_(p2, 2, cs)
    torch.cuda.synchronize()
    d = float((p1.float() - p2.float()).abs().max())
    assert d <= 3e-2, (rows, d)
    worst = max(worst, e, d)
    print(f"{it:3d} rows {rows:6d}: ffn max err {e:.4f}, proj fused vs unfused {d:.4f}", flush=True)
print("stress ok, worst", worst)
