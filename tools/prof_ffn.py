import sys
from pathlib import Path
import torch
sys.path.insert(0, "/root/repo")
from lgattn_pkg import pkg
M = pkg.matcher
dev = "cuda:0"
rows = 262144
x = torch.randn(rows, 512, device=dev).half()
w = (torch.randn(512, 512, device=dev) / 22.6).half()
bias = torch.randn(512, device=dev).half()
gamma = torch.randn(512, device=dev); beta = torch.randn(512, device=dev)
h1 = torch.empty(rows, 512, device=dev, dtype=torch.float16)
for _ in range(3):
    M.ffn1_ln_gelu(x, w, bias, gamma, beta, h1)
torch.cuda.synchronize()
print("ok")
