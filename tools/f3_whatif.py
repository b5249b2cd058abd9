"""Row f3 (one Q K^T for both cross-attention directions) as a what-if measurement: time the pair matcher's forward with the
product library and with a tuning build that drops the second direction's QK MMAs in the kv_batch_shift launches (wrong results,
timing only).  If the forward gains >= 5 % the transposed-softmax variant is worth building.

    python tools/f3_whatif.py build      # here
    python tools/f3_whatif.py run        # GPU box
"""
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "lightglue-with-flashattentionv2-tensorrt_b200" / "csrc"
LIB = CSRC / "build" / "variants" / "f3whatif" / "liblightglue_attention_plugin.so"

CHILD = r'''
import json, sys, torch
sys.path.insert(0, %r)
from lgattn_pkg import pkg
from oracle import matcher_oracle as mo
res = {}
for (pairs, kpts) in [(32, 2048), (64, 1024), (8, 4096)]:
    m = pkg.PairMatcher(seed=0, device="cuda:0")
    g = torch.Generator().manual_seed(kpts)
    kp = [torch.rand(pairs, kpts, 2, generator=g).cuda() * 2 - 1 for _ in range(2)]
    d = [torch.nn.functional.normalize(torch.randn(pairs, kpts, 256, generator=g), dim=-1).half().cuda() for _ in range(2)]
    for _ in range(3): m.match(kp[0], kp[1], d[0], d[1])
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): m.match(kp[0], kp[1], d[0], d[1])
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 10)
    res[f"{pairs}x{kpts}"] = round(best, 4)
print("RESULT " + json.dumps(res))
'''

if sys.argv[1] == "build":
    import importlib.util
    spec = importlib.util.spec_from_file_location("lgattn_build", CSRC / "build.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    print(mod.build(force=True, defs=["-DLG_ATTN_TUNING_BUILD", "-DLG_ATTN_WHATIF_F3"], out=LIB))
else:
    out = {}
    for rep in range(2):
        for tag, lib in (("product", None), ("f3_whatif", LIB)):
            env = dict(os.environ)
            if lib:
                env["LG_ATTN_LIB"] = str(lib)
            r = subprocess.run([sys.executable, "-c", CHILD % str(ROOT)], env=env, capture_output=True, text=True, timeout=600)
            line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
            out[f"{tag}_{rep}"] = json.loads(line[0][7:]) if line else {"error": r.stderr[-400:]}
            print(tag, rep, out[f"{tag}_{rep}"], flush=True)
    ks = out["product_0"].keys()
    for k in ks:
        p = min(out["product_0"][k], out["product_1"][k])
        w = min(out["f3_whatif_0"][k], out["f3_whatif_1"][k])
        print(f"forward ms (pairs x kpts {k}): product {p:.3f}  what-if {w:.3f}  gain {(p / w - 1) * 100:.1f} %")
