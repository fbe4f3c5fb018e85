"""Condenses an ncu launch list (--metrics gpu__time_duration.sum --csv) into per-kernel totals and shares.
usage: python tools/summarize_launches.py gpurun_out/launches.csv "<command that was profiled>" > profiles/rNN_launches_vK.csv"""
import csv
import re
import sys
from collections import defaultdict

path, what = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
rows = []
with open(path) as fh:
    lines = [ln for ln in fh if ln.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*$", "", r["Kernel Name"]).strip()
    rows.append((name, float(r["Metric Value"].replace(",", "")) * (1e-6 if r["Metric Unit"] == "ns" else 1e-3)))
tot = sum(t for _, t in rows)
agg = defaultdict(lambda: [0, 0.0])
for name, t in rows:
    agg[name][0] += 1
    agg[name][1] += t
print(f"# ncu launch list summary: {what}, gpu__time_duration.sum, --clock-control none")
print("# per-launch times under ncu are cold-cache and serialised: compare SHARES, not absolutes")
print(f"# total {tot:.3f} ms over {len(rows)} launches")
print("kernel,launches,total_ms,share")
for name, (cnt, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{name},{cnt},{t:.4f},{t / tot:.4f}")
