"""Summarise an `ncu --page source --csv` export: top instructions by stall samples + totals per stall reason.
usage: python tools/ncu_top.py file.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = rows[1]
ci = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = {h: 0 for h in stall_cols}
total = 0
for r in data:
    try:
        total += int(r[ci["# Samples"]])
    except ValueError:
        continue
    for h in stall_cols:
        tot[h] += int(r[ci[h]] or 0)
print("total samples", total)
print("by reason:", ", ".join(f"{h[6:]}={v} ({100*v/max(total,1):.1f}%)" for h, v in sorted(tot.items(), key=lambda x: -x[1]) if v))
data.sort(key=lambda r: -int(r[ci["# Samples"]] or 0))
for r in data[:n]:
    s = int(r[ci["# Samples"]] or 0)
    top = sorted(((int(r[ci[h]] or 0), h[6:]) for h in stall_cols), reverse=True)[:2]
    print(f"{s:6d} {100*s/max(total,1):5.1f}%  {r[ci['Source']].strip()[:90]:90s} {top}")
