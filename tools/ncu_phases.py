"""Stall-sample breakdown of an `ncu --page source --csv` export: top SASS lines and per-reason totals.
usage: python tools/ncu_phases.py gpurun_out/r1d_<layer>.source.csv [top_n]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 14
hdr = rows[1]
si, src = hdr.index("# Samples"), hdr.index("Source")
ie = hdr.index("Instructions Executed")
stalls = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data, reasons, ninst = [], collections.Counter(), 0
for k, r in enumerate(rows[2:]):
    try:
        n = int(r[si])
    except (ValueError, IndexError):
        continue
    ninst += int(r[ie] or 0)
    for i in stalls:
        reasons[hdr[i]] += int(r[i] or 0)
    top = max(((int(r[i] or 0), hdr[i]) for i in stalls))
    data.append((n, k, r[src][:80], top[1]))
tot = sum(d[0] for d in data)
print(f"kernel: {rows[0][1][:110]}")
print(f"samples {tot}, warp instructions {ninst}")
print("reasons:", ", ".join(f"{k[6:]} {100 * v / max(1, sum(reasons.values())):.0f}%" for k, v in reasons.most_common(7)))
for n, k, s, why in sorted(data, reverse=True)[:top_n]:
    print(f"{100 * n / tot:5.1f}%  line {k:5d}  {why[6:]:18s} {s}")
