"""Mnemonic histogram per tcgen05 kernel (attention, fused ffn-up, projection) of the built library (what proves a Blackwell-native kernel, B200_PROFILING.md):

    python tools/sass_histogram.py > profiles/r2_sass_histogram.txt
"""
import re
import subprocess
import sys
from collections import Counter, OrderedDict
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "lightglue-with-flashattentionv2-tensorrt_b200" / "csrc" / "build" / "liblightglue_attention_plugin.so"
WANT = ["UTCHMMA", "UTCBAR", "UTMALDG", "UTMASTG", "LDTM", "STTM", "MUFU.EX2", "MUFU.RCP", "FFMA2", "FADD2", "FFMA", "FMNMX3", "F2FP", "ELECT", "BRA.U.ANY",
        "SYNCS", "UCGABAR", "HMMA", "HGMMA", "LDL", "STL"]

sass = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True).stdout
kernels = OrderedDict()
cur = None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        kernels[cur] = Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,5}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1)
        kernels[cur]["_total"] += 1
        for w in WANT:
            if op == w or op.startswith(w + "."):
                kernels[cur][w] += 1
                break
print(f"# {LIB.relative_to(ROOT)}: SASS mnemonic counts per kernel (cuobjdump -sass); UTCHMMA = tcgen05.mma, UTMALDG = TMA tensor load,")
print("# LDTM / STTM = tcgen05.ld / st, UTCBAR = tcgen05.commit, BRA.U.ANY = per-instruction waterfall loops (0 since the service warps")
print("# issue through elect.sync), HMMA / HGMMA = legacy mma.sync / wgmma (must be 0), LDL / STL = spills")
print("kernel".ljust(78) + " ".join(w.rjust(9) for w in WANT) + "   total")
for name, c in kernels.items():
    if not any(k in name for k in ("lg_attn_hd64_fwd", "ffn1_ln_gelu", "proj_rotary_kernel")) and "--all" not in sys.argv:
        continue
    short = re.sub(r"\(.*", "", name).replace("lgattn::", "")
    print(short[:76].ljust(78) + " ".join(str(c[w]).rjust(9) for w in WANT) + f"   {c['_total']}")
