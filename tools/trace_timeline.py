"""GPU box, trace build only (LG_ATTN_LIB=…/variants/trace/…): per-launch timeline of a CUDA-graph
chain of B1 H4 N=M problems, from the kernel's %globaltimer stamps (CTAs 0-3 = first cluster)."""
import ctypes, os, sys
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from lgattn_pkg import pkg

lib = pkg.load()
n = int(os.environ.get("N", "1024"))
calls = 32
q = torch.randn(calls, 1, 4, n, 64, device="cuda").half(); k = torch.randn_like(q); v = torch.randn_like(q); o = torch.empty_like(q)
def go(s):
    for i in range(calls):
        assert lib.lg_attn_hd64_launch(q[i].data_ptr(), k[i].data_ptr(), v[i].data_ptr(), o[i].data_ptr(), 1, 4, n, n, 0, None, 0, s) == 0
st = torch.cuda.Stream(); g = torch.cuda.CUDAGraph()
NEV = 14
buf = np.zeros((64, 4, NEV), dtype=np.uint64)
with torch.cuda.stream(st):
    go(st.cuda_stream); st.synchronize()
    lib.lg_attn_debug_read_trace(buf.ctypes.data_as(ctypes.c_void_p), 1)
    with torch.cuda.graph(g, stream=st): go(torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
for _ in range(3): g.replay()
torch.cuda.synchronize()
lib.lg_attn_debug_read_trace(buf.ctypes.data_as(ctypes.c_void_p), 0)
order = [12, 13, 1, 2, 3, 4, 5, 10, 11, 6, 7, 8, 9]
names = {0: "entry", 1: "setup", 2: "Qfull", 3: "S0full", 4: "loopend", 5: "PVdone", 6: "pushed", 7: "landed", 8: "stored", 9: "exit",
         10: "staged", 11: "cwait1", 12: "pdlwait", 13: "tmaiss"}
t = buf.astype(np.int64)
print("launch/cta: gap | " + " ".join(f"{names[x]:>7s}" for x in order) + "   (ns since CTA 0 entry)")
for i in range(16, 20):
    e0 = t[i, 0, 0]
    gap = e0 - t[i - 1, :, 9].max()
    for c in range(4 if int(os.environ.get("ALLCTAS", "0")) else 1):
        e = t[i, c]
        print(f"{i:3d}/{c}: {gap:5d} | " + " ".join(f"{int(e[x] - e0) if e[x] else -1:7d}" for x in order) + f"  entry+{int(e[0]-e0)}")
d = np.diff(t[8:32, 0, 0])
print("entry-to-entry ns: mean", d.mean(), "min", d.min(), "max", d.max())
if hasattr(lib, "lg_attn_debug_max_clusters"):
    print("max active clusters (splits, extra smem):", {(s_, ex): lib.lg_attn_debug_max_clusters(s_, ex) for s_ in (2, 4, 8) for ex in (0, 1)})
