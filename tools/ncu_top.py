"""top stall sites of an ncu source-page CSV (ncu -i X.ncu-rep --page source --csv > X.csv): python tools/ncu_top.py X.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]
c, si, ie = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
data = [r for r in rows[hi + 1:] if len(r) > c]
tot = sum(float(r[c] or 0) for r in data)
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
print(f"total samples {tot:.0f}, instructions {len(data)}")
for i, r in sorted(enumerate(data), key=lambda x: -float(x[1][c] or 0))[:n]:
    print(f"{float(r[c]) / tot * 100:5.1f}%  #{i:5d} exec={r[ie]:>8s}  {r[si].strip()[:100]}")
