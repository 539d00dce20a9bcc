"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list into a markdown table
(share of total, total us, launches, kernel).  usage: python scripts/launch_summary.py in.csv "title" "command" > out.md"""
import csv
import sys
from collections import defaultdict

path, title, cmd = sys.argv[1], sys.argv[2], sys.argv[3]
rows = []
with open(path, newline="") as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
tot = defaultdict(float)
cnt = defaultdict(int)
n = 0
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    u = r.get("Metric Unit", "ns")
    v_us = v / 1e3 if u in ("ns", "nsecond") else (v if u in ("us", "usecond") else v * 1e3)
    k = r["Kernel Name"]
    tot[k] += v_us
    cnt[k] += 1
    n += 1
total = sum(tot.values())
print(f"# {title}\n\nCommand: `{cmd}`\n(cold-cache serialised per-launch times: compare SHARES, not absolutes)\n")
print(f"total {total / 1e3:.1f} ms over {n} launches\n\n| share | total us | launches | kernel |\n|---|---|---|---|")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:60]:
    print(f"| {100 * v / total:.2f}% | {v:.1f} | {cnt[k]} | `{k[:110]}` |")
