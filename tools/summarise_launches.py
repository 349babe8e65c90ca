"""ncu launch list (--metrics gpu__time_duration.sum --csv) -> per-kernel summary.
python tools/summarise_launches.py gpurun_out/launches.csv "header line" > profiles/rX_launches_summary.txt"""
import csv, re, sys
from collections import defaultdict

rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") == "gpu__time_duration.sum":
        name = re.sub(r"\(.*$", "", r["Kernel Name"]).replace("pyn::", "")
        ns = float(r["Metric Value"].replace(",", "")) * {"ns": 1.0, "us": 1e3, "ms": 1e6}.get(r["Metric Unit"], 1.0)
        rows.append((name, ns))
tot = sum(ns for _, ns in rows)
agg = defaultdict(lambda: [0, 0.0])
for name, ns in rows:
    agg[name][0] += 1
    agg[name][1] += ns
for h in sys.argv[2:]:
    print("# " + h)
print(f"# {len(rows)} launches, {tot / 1e6:.3f} ms of kernel time")
for name, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{name:50s} n={n:6d} total={ns / 1e6:9.3f} ms avg={ns / n / 1e3:9.2f} us share={ns / tot:.4f}")
