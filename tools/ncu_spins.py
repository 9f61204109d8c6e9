"""who waits on what in the MMQ kernel: spin counts of every mbarrier try_wait loop (ncu source-page CSV)"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]; c = hdr.index("# Samples"); si = hdr.index("Source"); ie = hdr.index("Instructions Executed")
data = rows[hi + 1:]
tot = sum(float(r[c] or 0) for r in data)
for i, r in enumerate(data):
    if "TRYWAIT" in r[si] and int(r[ie] or 0) > 0:
        nxt = data[i + 1]
        print(f"#{i:5d} exec={int(r[ie]):9d} samples={float(r[c]) + float(nxt[c] or 0):6.0f} ({(float(r[c]) + float(nxt[c] or 0)) / tot * 100:4.1f}%)  {r[si].strip()[32:]}")
