"""Summarise an `ncu --page source --csv --print-source sass` export: stall samples per code region
(regions are delimited by synchronisation / TMEM / MMA instructions)."""
import csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}; data = rows[2:]
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
acc = 0; accst = {h: 0 for h in stalls}; n = 0
tot = sum(int(r[ix['# Samples']] or 0) for r in data)
print('total samples', tot)
for r in data:
    s = int(r[ix['# Samples']] or 0); src = r[ix['Source']]
    acc += s; n += 1
    for h in stalls: accst[h] += int(r[ix[h]] or 0)
    if re.search(r'SYNCS|LDTM|STTM|UTCHMMA|UTMALDG|BAR|EXIT|UTCBAR|USETMAXREG|VOTE', src):
        if acc >= int(sys.argv[2]) if len(sys.argv) > 2 else 50:
            top = sorted(accst.items(), key=lambda kv: -kv[1])[:4]
            print(r[ix['Address']][-5:], f"samples={acc:5d} ({100*acc/tot:4.1f}%) instrs={n:4d}", src[:60], r[ix['Instructions Executed']], [(k[6:], v) for k, v in top if v])
        acc = 0; n = 0; accst = {h: 0 for h in stalls}
