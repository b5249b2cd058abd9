"""GPU box: chained latency (us / problem in a CUDA-graph chain) of the B1 H4 N=M=1024 op for each io mode, for library variants.

    python tools/lat_probe.py [tag ...]      (tag 'main' = product build; others = csrc/build/variants/<tag>)
"""
import ctypes
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import tools.ab_bench as ab  # noqa: E402
import tools.perf_compare as pc  # noqa: E402

tags = sys.argv[1:] or ["main"]
libs = {t: ab.load(t) for t in tags}
calls, n, m = 128, 1024, 1024
for name, mode, tin, tout in (("f16->f16", 0, torch.float16, torch.float16), ("f16->f32", 2, torch.float16, torch.float32), ("f32->f32", 1, torch.float32, torch.float32)):
    q = torch.randn(calls, 1, 4, n, 64, device="cuda").to(tin)
    k = torch.randn_like(q)
    v = torch.randn_like(q)
    o = torch.empty(calls, 1, 4, n, 64, device="cuda", dtype=tout)
    for rep in range(2):
        for t, lib in libs.items():
            ws = getattr(lib, "lg_attn_hd64_workspace_bytes")
            ws.restype = ctypes.c_size_t
            ws.argtypes = [ctypes.c_int] * 5
            nb = ws(1, 4, n, m, mode)
            wsb = torch.empty(max(nb, 16), dtype=torch.uint8, device="cuda")

            def ours(i, s, lib=lib):
                assert lib.lg_attn_hd64_launch(q[i].data_ptr(), k[i].data_ptr(), v[i].data_ptr(), o[i].data_ptr(), 1, 4, n, m, mode,
                                               wsb.data_ptr() if nb else None, nb, s) == 0

            print(f"{name:9s} {t:8s} chained {pc.time_graph_chain(ours, calls):6.2f} us/problem", flush=True)
