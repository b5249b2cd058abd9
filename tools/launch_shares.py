"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list into per-kernel shares.

    python tools/launch_shares.py gpurun_out/launches.csv "header comment" > profiles/rN_launches_x.csv
"""
import csv
import re
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = OrderedDict()
for r in rows[1:]:
    if len(r) <= vi:
        continue
    name = re.sub(r"\(.*", "", r[ki])[:90]
    ns = float(r[vi].replace(",", "")) * {"ns": 1.0, "us": 1e3, "ms": 1e6}.get(r[ui], 1.0)
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += ns
tot = sum(a[1] for a in agg.values())
if len(sys.argv) > 2:
    print("# " + sys.argv[2])
print("# per-launch times under ncu are cold-cache and serialised: compare SHARES, not absolutes")
print("kernel,launches,total_ns,mean_ns,share")
for name, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f'"{name}",{n},{ns:.0f},{ns / n:.0f},{ns / tot:.4f}')
print(f'"TOTAL",{sum(a[0] for a in agg.values())},{tot:.0f},,1.0')
