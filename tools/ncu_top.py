"""Top stalled SASS lines of an ncu source-page CSV: python tools/ncu_top.py file.csv [n]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
h = rows[1]
ci, si, ii = h.index("Warp Stall Sampling (All Samples)"), h.index("Source"), h.index("Instructions Executed")
stall_cols = [i for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
data = []
for r in rows[2:]:
    try:
        st = sorted(((float(r[i]), h[i][6:]) for i in stall_cols), reverse=True)[:2]
        data.append((float(r[ci]), r[si].strip()[:90], r[ii], st))
    except (ValueError, IndexError):
        pass
tot = sum(d[0] for d in data) or 1
for d in sorted(data, reverse=True)[:n]:
    print(f"{100*d[0]/tot:5.1f}% {d[2]:>9} {d[1]:<90} {d[3][0][1]}:{int(d[3][0][0])} {d[3][1][1]}:{int(d[3][1][0])}")
