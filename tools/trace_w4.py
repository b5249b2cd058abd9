"""GPU box, trace build only (LG_ATTN_LIB=…/variants/w4trace/…): per-tile phase durations of the
throughput kernel's softmax warp 0 (clock64), for CTAs resident in the steady state of a B64 N2048 launch."""
import ctypes, os, sys
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from lgattn_pkg import pkg

lib = pkg.load()
b, n = int(os.environ.get("B", "64")), int(os.environ.get("N", "2048"))
q = torch.randn(b, 4, n, 64, device="cuda").half(); k = torch.randn_like(q); v = torch.randn_like(q)
for _ in range(3):
    pkg.attention(q, k, v)
torch.cuda.synchronize()
NC, NT, NE = 592, 16, 8
buf = np.zeros((NC, NT, NE), dtype=np.int64); smid = np.zeros(NC, dtype=np.uint32)
assert lib.lg_attn_debug_read_w4_trace(buf.ctypes.data_as(ctypes.c_void_p), smid.ctypes.data_as(ctypes.c_void_p)) == 0
names = ["waitS", "ldS", "max", "exp", "waitPV", "stP+arrive", "loop-gap"]
# CTAs 296..591 are second-wave (steady state: both co-resident CTAs are mid-flight)
sel = buf[296:592]
d = np.diff(sel[:, :, :7], axis=2).astype(np.float64)            # [cta, tile, 6 phases]
gap = (sel[:, 1:, 0] - sel[:, :-1, 6]).astype(np.float64)         # end of tile j -> start of tile j+1
per_tile = (sel[:, 1:, 0] - sel[:, :-1, 0]).astype(np.float64)
print("phase means over tiles 2..14 (cycles):")
for i, nm in enumerate(names[:6]):
    print(f"  {nm:12s} mean {d[:, 2:15, i].mean():8.1f}  p50 {np.median(d[:, 2:15, i]):8.1f}  p90 {np.percentile(d[:, 2:15, i], 90):8.1f}")
print(f"  {'loop-gap':12s} mean {gap[:, 2:14].mean():8.1f}")
print(f"  tile period  mean {per_tile[:, 2:14].mean():8.1f}  p50 {np.median(per_tile[:, 2:14]):8.1f}")
# phase offset between the two CTAs sharing an SM
by_sm = {}
for c in range(296, 592):
    by_sm.setdefault(int(smid[c]), []).append(c)
offs = []
for sm, cs in by_sm.items():
    if len(cs) >= 2:
        a, bb = cs[0], cs[1]
        t = (buf[bb, 5, 3] - buf[a, 5, 3]) % max(1.0, per_tile.mean())
        offs.append(t / per_tile.mean())
print("co-resident CTA phase offset (fraction of tile period): ", np.round(np.percentile(offs, [10, 50, 90]), 2) if offs else "n/a")
c0 = 300
print("one CTA, tiles 3..6, stamps relative to tile start:", (buf[c0, 3:7, :7] - buf[c0, 3:7, :1]).tolist())

# CTA-level stamps (slot 7 of tile rows 0..5)
cta = buf[:, :6, 7].astype(np.float64)
ok = cta[:, 0] > 0
setup = cta[ok, 1] - cta[ok, 0]; loop = cta[ok, 2] - cta[ok, 1]; epi = cta[ok, 3] - cta[ok, 2]; life = cta[ok, 3] - cta[ok, 0]
first = buf[ok, 0, 1] - cta[ok, 1]
print(f"CTA clocks: set-up {np.median(setup):.0f}  first S wait {np.median(first):.0f}  tile loop {np.median(loop):.0f}  epilogue {np.median(epi):.0f}  lifetime {np.median(life):.0f}")
for lo, hi, tag in ((0, 296, "wave 1"), (296, 592, "wave 2")):
    l = (cta[lo:hi, 3] - cta[lo:hi, 0]); print(f"  {tag}: lifetime p50 {np.median(l):.0f} p10 {np.percentile(l,10):.0f} p90 {np.percentile(l,90):.0f}; loop p50 {np.median(cta[lo:hi,2]-cta[lo:hi,1]):.0f}; ns p50 {np.median(cta[lo:hi,5]-cta[lo:hi,4]):.0f}")
# successor gap on the same SM: wave-2 CTA start minus the end of the wave-1 CTA it replaced (globaltimer ns)
gaps = []
for sm, cs in by_sm.items():
    firsts = [c for c in range(296) if int(smid[c]) == sm]
    if not firsts: continue
    ends = sorted(buf[c, 5, 7] for c in firsts)
    starts = sorted(buf[c, 4, 7] for c in cs)
    for e, st in zip(ends, starts): gaps.append(st - e)
if gaps: print("same-SM successor gap ns (p10/p50/p90):", np.percentile(gaps, [10, 50, 90]))
