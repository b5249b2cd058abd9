"""Profiling driver: ONE eager forward of the pair matcher (8 pairs x 2048 keypoints, 9 layers) after a
warm-up forward — run under `ncu --metrics gpu__time_duration.sum` for the per-kernel launch list."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from lgattn_pkg import pkg  # noqa: E402
from tools.pairs_sweep import feats  # noqa: E402

dev = torch.device("cuda:0")
b, n = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (8, 2048)
m = pkg.PairMatcher(seed=0, device=dev, use_graph=False)
f = feats(b, n, dev)
for _ in range(2):
    r = m.match(*f)
torch.cuda.synchronize()
print("ok launches per forward:", r["launches"])
