"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel:
python tools/launch_summary.py gpurun_out/rNN_launches.csv > profiles/rNN_launches_summary.txt"""
import csv, re, sys
from collections import defaultdict

rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r["Metric Name"] == "gpu__time_duration.sum":
        rows.append((r["Kernel Name"], r["Grid Size"], r["Block Size"], float(r["Metric Value"].replace(",", ""))))

def short(name):
    name = re.sub(r"^void ", "", name).replace("(anonymous namespace)::", "").replace("<unnamed>::", "")
    depth = 0
    for i, ch in enumerate(name):                     # cut the argument list, keep template arguments
        depth += ch == "<"
        depth -= ch == ">"
        if ch == "(" and depth == 0:
            name = name[:i]
            break
    return name[:90]

agg = defaultdict(lambda: [0, 0.0, None])
for name, grid, block, ns in rows:
    a = agg[short(name)]
    a[0] += 1
    a[1] += ns
    a[2] = f"{grid} x {block}"
total = sum(a[1] for a in agg.values())
ours = sum(a[1] for k, a in agg.items() if k.startswith("ocppo::"))
print(f"# {len(rows)} launches, {total / 1e6:.3f} ms summed kernel time (cold-cache, serialised by ncu); "
      f"ocppo:: kernels {100 * ours / total:.1f} % of it")
print(f"{'kernel':<92}{'launches':>9}{'total ms':>11}{'avg us':>10}{'share':>8}  last grid x block")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:<92}{a[0]:>9}{a[1] / 1e6:>11.3f}{a[1] / a[0] / 1e3:>10.1f}{100 * a[1] / total:>7.1f}%  {a[2]}")
