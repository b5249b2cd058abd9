"""Tuning aid: build kernel variants (extra -D flags) here, time them on the GPU box.

    python tools/variant_sweep.py build          # in the build container: compiles every variant
    python tools/variant_sweep.py run            # on the GPU box: correctness + timing per variant
"""
import importlib.util
import json
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "lightglue-with-flashattentionv2-tensorrt_b200" / "csrc"
VDIR = CSRC / "build" / "variants"

VARIANTS = {
    # tag: defs (both kernels of the library share the softmax arithmetic knobs)
    "default": [],
    "p0": ["-DLG_ATTN_POLY_OF_8=0"],
    "p1": ["-DLG_ATTN_POLY_OF_8=1"],
    "p3": ["-DLG_ATTN_POLY_OF_8=3"],
    "p2_scalar_scale": ["-DLG_ATTN_PACKED_SCALE=0"],
    "p2_scalar_sum": ["-DLG_ATTN_PACKED_SUM=0"],
    "spin32": ["-DLG_ATTN_SPIN_NS=32"],
    "spin128": ["-DLG_ATTN_SPIN_NS=128"],
    "tau6": ["-DLG_ATTN_TAU=6.0f"],
}


def build_all(tags):
    spec = importlib.util.spec_from_file_location("lgattn_build", CSRC / "build.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)

    def one(tag):
        out = VDIR / tag / "liblightglue_attention_plugin.so"
        mod.build(force=True, defs=VARIANTS[tag], out=out)
        return tag

    with ThreadPoolExecutor(6) as ex:
        for t in ex.map(one, tags):
            print("built", t, flush=True)


CHILD = r'''
import json, sys, torch
sys.path.insert(0, %r)
from lgattn_pkg import pkg
from oracle import attention_oracle as orc
import numpy as np
res = {}
q, k, v = orc.make_qkv(1, 1, 4, 777, 1000)
ref = orc.attention_np(q, k, v)
tq, tk, tv = (torch.from_numpy(x).cuda() for x in (q, k, v))
res["err16"] = float(orc.max_abs_err(pkg.attention(tq, tk, tv).float().cpu().numpy(), ref))
res["err32"] = float(orc.rel_err(pkg.attention(tq, tk, tv, out_dtype=torch.float32).cpu().numpy(), ref))
def timeit(b, n, iters=20):
    q = torch.randn(b, 4, n, 64, device="cuda").half(); k = torch.randn_like(q); v = torch.randn_like(q); o = torch.empty_like(q)
    for _ in range(3): pkg.attention(q, k, v, out=o)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): pkg.attention(q, k, v, out=o)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters
    return us, 4.0 * b * 4 * n * n * 64 / us / 1e6
res["B64_N2048_us"], res["B64_N2048_tflops"] = timeit(64, 2048)
res["B16_N4096_us"], res["B16_N4096_tflops"] = timeit(16, 4096)
# latency: graph chain of 64 independent B1 N1024 problems
calls = 64
q = torch.randn(calls, 1, 4, 1024, 64, device="cuda").half(); k = torch.randn_like(q); v = torch.randn_like(q); o = torch.empty_like(q)
lib = pkg.load()
st = torch.cuda.Stream(); g = torch.cuda.CUDAGraph()
def go(s):
    for i in range(calls):
        assert lib.lg_attn_hd64_launch(q[i].data_ptr(), k[i].data_ptr(), v[i].data_ptr(), o[i].data_ptr(), 1, 4, 1024, 1024, 0, None, 0, s) == 0
with torch.cuda.stream(st):
    go(st.cuda_stream); st.synchronize()
    with torch.cuda.graph(g, stream=st): go(torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
for _ in range(5): g.replay()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): g.replay()
e1.record(); torch.cuda.synchronize()
res["B1_N1024_us"] = e0.elapsed_time(e1) * 1e3 / (20 * calls)
print("RESULT " + json.dumps(res))
'''


def run_all(tags):
    out = {}
    for tag in tags:
        lib = VDIR / tag / "liblightglue_attention_plugin.so"
        if not lib.exists():
            continue
        import time
        time.sleep(3.0)  # let power / clocks recover: every variant is timed as a burst from idle
        env = dict(os.environ, LG_ATTN_LIB=str(lib))
        r = subprocess.run([sys.executable, "-c", CHILD % str(ROOT)], env=env, capture_output=True, text=True, timeout=300)
        line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
        out[tag] = json.loads(line[0][7:]) if line else {"error": (r.stderr or r.stdout)[-400:]}
        print(tag, json.dumps(out[tag]), flush=True)
    (ROOT / "gpurun_out").mkdir(exist_ok=True)
    (ROOT / "gpurun_out" / "variant_sweep.json").write_text(json.dumps(out, indent=1))


if __name__ == "__main__":
    tags = sys.argv[2:] or list(VARIANTS)
    (build_all if sys.argv[1] == "build" else run_all)(tags)
