"""compute-sanitizer target (one tool per gpurun call): a handful of ragged attention launches through
both kernels and both entries, plus one small matcher forward."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from lgattn_pkg import pkg
from oracle import matcher_oracle as mo

dev = "cuda:0"
for (b, n, m) in [(1, 1, 1), (1, 129, 65), (1, 1024, 1000), (2, 300, 333), (80, 130, 257)]:
    q = torch.randn(b, 4, n, 64, device=dev).half(); k = torch.randn(b, 4, m, 64, device=dev).half(); v = torch.randn_like(k)
    pkg.attention(q, k, v)
    pkg.attention(q, k, v, out_dtype=torch.float32)
    pkg.attention(q.float(), k.float(), v.float())
sd = pkg.reference_init_state_dict(1, 2)
for (n0, n1) in [(70, 70), (33, 129)]:
    f = mo.synth_pair(5, n0, n1)
    pkg.PairMatcher(sd, n_layers=2, device=dev, use_graph=False).match(*(t.to(dev) for t in f), want_scores=True)
torch.cuda.synchronize()
print("sanitize target ok")
