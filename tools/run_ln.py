"""GPU box: a few LN+GELU launches at the pair-matcher's size (the command ncu wraps)."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from lgattn_pkg import pkg
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
h = torch.randn(rows, 512, device="cuda").half()
g = torch.ones(512, device="cuda"); b = torch.zeros(512, device="cuda")
for _ in range(6):
    pkg.matcher.layernorm_gelu_(h, g, b)
torch.cuda.synchronize()
print("ok")
