"""Hottest SASS instructions of a capture by warp-stall samples, with their dominant stall reason:
   python scripts/ncu_hot.py file.ncu-rep [top_n]"""
import csv
import io
import subprocess
import sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
head = rows[hi]
col = {n: i for i, n in enumerate(head)}
stalls = [n for n in head if n.startswith("stall_") and "Not Issued" not in n]
body = [r for r in rows[hi + 1:] if len(r) == len(head)]
tot = sum(int(r[col["# Samples"]] or 0) for r in body)
print("total samples", tot, "instructions", len(body))
agg = {}
for r in body:
    for s in stalls:
        agg[s] = agg.get(s, 0) + int(r[col[s]] or 0)
print("stall totals:", ", ".join(f"{k[6:]}={v}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
order = sorted(range(len(body)), key=lambda i: -int(body[i][col["# Samples"]] or 0))[:top]
for i in sorted(order):
    r = body[i]
    n = int(r[col["# Samples"]] or 0)
    st = sorted(((int(r[col[s]] or 0), s[6:]) for s in stalls), reverse=True)[:2]
    print(f"{i:5d} {100 * n / tot:5.1f}%  exec={r[col['Instructions Executed']]:>9s}  {r[col['Source']].strip()[:70]:70s} {st}")
