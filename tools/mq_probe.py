"""GPU box: correctness + burst TFLOP/s of every full-grid kernel the library carries, selected through
lg_attn_debug_set_throughput_kernel (0 = w4, 1 = mq 4 Q tiles, 2 = mq 2 Q tiles; -2 = torch SDPA for reference).
Each mode runs in its own subprocess under a timeout so that a hung variant cannot take the call down.

    python tools/mq_probe.py [--modes 0,2,3] [--out gpurun_out/mq_probe.json] [--lib path/to/variant.so]
"""
import argparse
import json
import os
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent

CHILD = r'''
import json, sys, torch
sys.path.insert(0, %(root)r)
from lgattn_pkg import pkg
mode = %(mode)d
lib = pkg.load()
if mode >= 0:
    assert lib.lg_attn_debug_set_throughput_kernel(mode) == 0
res = {"mode": mode}
torch.manual_seed(1)
def ref_slices(q, k, v, idx):
    out = []
    for (b, h) in idx:
        s = (q[b, h].float() @ k[b, h].float().T) * 0.125
        out.append(torch.softmax(s, -1) @ v[b, h].float())
    return torch.stack(out)
def check(b, n, m, scale=1.0, peaked=False):
    q = (torch.randn(b, 4, n, 64, device="cuda") * scale).half()
    k = (torch.randn(b, 4, m, 64, device="cuda") * scale).half()
    v = torch.randn(b, 4, m, 64, device="cuda").half()
    if peaked:  # growing row max along the KV axis: exercises the rescale / lazy-max slow path
        k = (k.float() * torch.linspace(0.2, 3.0, m, device="cuda")[None, None, :, None]).half()
    idx = [(0, 0), (b - 1, 3), (b // 2, 1)]
    ref = ref_slices(q, k, v, idx)
    o16 = pkg.attention(q, k, v)
    o32 = pkg.attention(q, k, v, out_dtype=torch.float32)
    g16 = torch.stack([o16[bi, hi].float() for (bi, hi) in idx])
    g32 = torch.stack([o32[bi, hi] for (bi, hi) in idx])
    e16 = float((g16 - ref).abs().max())
    e32 = float((g32 - ref).abs().max() / ref.abs().max())
    return e16, e32
errs = {}
for (b, n, m, sc, pk) in [(160, 128, 128, 1.0, False), (80, 777, 1000, 1.0, False), (40, 300, 2047, 1.0, False), (150, 512, 64, 1.0, False),
                          (150, 500, 1, 1.0, False), (64, 1024, 1024, 2.0, True), (40, 640, 1500, 1.5, True), (16, 4096, 4096, 1.0, False)]:
    e = check(b, n, m, sc, pk)
    errs[f"B{b}_N{n}_M{m}_s{sc}_{int(pk)}"] = ["%%.2e" %% x for x in e]
    res["ok"] = res.get("ok", True) and e[0] <= 2e-3 and e[1] <= 1e-3
res["errs"] = errs
def timeit(b, n, iters=20, f32=False):
    q = torch.randn(b, 4, n, 64, device="cuda").half(); k = torch.randn_like(q); v = torch.randn_like(q)
    o = torch.empty(b, 4, n, 64, device="cuda", dtype=torch.float32 if f32 else torch.float16)
    fn = (lambda: torch.nn.functional.scaled_dot_product_attention(q, k, v)) if mode == -2 else (lambda: pkg.attention(q, k, v, out_dtype=o.dtype, out=o))
    for _ in range(5): fn()
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters): fn()
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / iters)
    return round(4.0 * b * 4 * n * n * 64 / (best * 1e-3) / 1e12, 1)
res["tflops"] = {"B64_N2048": timeit(64, 2048), "B16_N4096": timeit(16, 4096), "B64_N4096": timeit(64, 4096, iters=8), "B128_N1024": timeit(128, 1024),
                 "B256_N512": timeit(256, 512), "B64_N2048_f32": timeit(64, 2048, f32=True)}
print("RESULT " + json.dumps(res))
'''


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--modes", default="-2,0,1,2")
    ap.add_argument("--out", default="gpurun_out/mq_probe.json")
    ap.add_argument("--lib", default=None)
    ap.add_argument("--tag", default="")
    args = ap.parse_args()
    out = {}
    env = dict(os.environ)
    if args.lib:
        env["LG_ATTN_LIB"] = args.lib
    for mode in [int(x) for x in args.modes.split(",")]:
        time.sleep(2.0)
        try:
            r = subprocess.run([sys.executable, "-c", CHILD % {"root": str(ROOT), "mode": mode}], env=env, capture_output=True,
                               text=True, timeout=240)
            line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
            out[mode] = json.loads(line[0][7:]) if line else {"error": (r.stderr or r.stdout)[-600:]}
        except subprocess.TimeoutExpired:
            out[mode] = {"error": "timeout (hang?)"}
        print(args.tag, mode, json.dumps(out[mode]), flush=True)
    p = ROOT / args.out
    p.parent.mkdir(exist_ok=True)
    prev = json.loads(p.read_text()) if p.exists() else {}
    prev[args.tag or "default"] = out
    p.write_text(json.dumps(prev, indent=1))


if __name__ == "__main__":
    main()
