"""Phase trace of the multi-Q-tile kernel (tuning build with -DLG_ATTN_MQ_TRACE).

    python tools/trace_mq.py build                 # here: csrc/build/variants/mqtrace/
    python tools/trace_mq.py run MODE [B N]        # GPU box: prints per-step timelines (clk) of CTA 0
"""
import ctypes
import importlib.util
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "lightglue-with-flashattentionv2-tensorrt_b200" / "csrc"
LIB = CSRC / "build" / "variants" / "mqtrace" / "liblightglue_attention_plugin.so"

if sys.argv[1] == "build":
    spec = importlib.util.spec_from_file_location("lgattn_build", CSRC / "build.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    print(mod.build(force=True, defs=["-DLG_ATTN_MQ_TRACE"] + sys.argv[2:], out=LIB))
    sys.exit(0)

os.environ["LG_ATTN_LIB"] = str(LIB)
import numpy as np  # noqa: E402
import torch  # noqa: E402

sys.path.insert(0, str(ROOT))
from lgattn_pkg import pkg  # noqa: E402

mode = int(sys.argv[2])
b, n = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (64, 2048)
lib = pkg.load()
assert lib.lg_attn_debug_set_throughput_kernel(mode) == 0
q = torch.randn(b, 4, n, 64, device="cuda").half()
k = torch.randn_like(q)
v = torch.randn_like(q)
for _ in range(3):
    o = pkg.attention(q, k, v)
torch.cuda.synchronize()
CT, ST, TL, EV = 8, 64, 4, 8
buf = np.zeros((CT, ST, TL, EV), dtype=np.int64)
lib.lg_attn_debug_read_mq_trace.argtypes = [ctypes.c_void_p]
assert lib.lg_attn_debug_read_mq_trace(buf.ctypes.data) == 0
names = ["pre_wait", "s_full", "S_in_regs", "exp_done", "p_full_arr", "mma_saw_p", "mma_issued"]
for cta in (0, 5):
    t = buf[cta]
    nt = int((t[8, :, 1] != 0).sum())
    nsteps = int((t[:, 0, 1] != 0).sum())
    print(f"--- CTA {cta}: tiles {nt}, traced steps {nsteps}")
    base = t[8, 0, 1]
    for step in range(8, min(14, nsteps)):
        for tile in range(nt):
            ev = t[step, tile, :7] - base
            print(f"step {step:2d} tile {tile}: " + "  ".join(f"{nm}={int(x):6d}" for nm, x in zip(names, ev)))
    per = np.diff(t[8:nsteps, :nt, 1], axis=0)
    print("tile period (s_full to s_full) mean per tile:", per.mean(axis=0).round(0), " overall", per.mean().round(0))
    d = t[8:nsteps, :nt]
    print("mean clk: wait_s", (d[..., 1] - d[..., 0]).mean().round(0), " ldS", (d[..., 2] - d[..., 1]).mean().round(0), " exp", (d[..., 3] - d[..., 2]).mean().round(0),
          " st+arrive", (d[..., 4] - d[..., 3]).mean().round(0), " p_arrive->mma_saw", (d[..., 5] - d[..., 4]).mean().round(0),
          " mma_issue", (d[..., 6] - d[..., 5]).mean().round(0), " mma_issued->next s_full", (d[1:, :, 1] - d[:-1, :, 6]).mean().round(0))

# item boundaries: per-step s_full stamps of tile 0 over the whole trace (a step that starts a new item shows the boundary bubble)
if os.environ.get("MQ_TRACE_STEPS"):
    t = buf[0]
    nsteps = int((t[:, 0, 1] != 0).sum())
    prev = None
    for step in range(nsteps):
        ev = t[step, 0, :7]
        line = f"step {step:2d}: s_full {int(ev[1] - t[0, 0, 1]):7d}"
        if prev is not None:
            line += f"  (+{int(ev[1] - prev):5d})  wait {int(ev[1] - ev[0]):5d}  ld {int(ev[2] - ev[1]):4d}  exp {int(ev[3] - ev[2]):5d}  st {int(ev[4] - ev[3]):4d}  p->mma {int(ev[5] - ev[4]):5d}  issue {int(ev[6] - ev[5]):5d}"
        prev = ev[1]
        print(line)
