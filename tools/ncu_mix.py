"""Executed warp-instruction mix from an `ncu --page source --csv --print-source sass` export,
normalised per unit (argv[2] = number of units, e.g. CTA-tiles)."""
import csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}; data = rows[2:]
units = float(sys.argv[2])
tot = 0; per = {}
for r in data:
    n = int(r[ix['Instructions Executed']] or 0)
    src = re.sub(r'@!?U?P\d+\s+', '', r[ix['Source']].strip())
    op = src.split()[0].split('.')[0] if src else '?'
    per[op] = per.get(op, 0) + n; tot += n
print('total warp-instrs', tot, 'per unit', round(tot / units, 1))
print(' '.join(f"{k}:{v/units:.1f}" for k, v in sorted(per.items(), key=lambda kv: -kv[1])[:28]))
