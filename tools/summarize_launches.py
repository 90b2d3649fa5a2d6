"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import csv, re, sys
from collections import defaultdict
rows = []
with open(sys.argv[1]) as fh:
    for line in fh:
        if line.startswith('"ID"'):
            break
    rd = csv.reader(fh)
    for r in rd:
        if len(r) >= 15:
            rows.append((r[4], r[7], r[8], float(r[14].replace(",", ""))))
tot = sum(r[3] for r in rows)
agg = defaultdict(lambda: [0, 0.0, 0.0])
for name, blk, grid, ns in rows:
    m = re.search(r"(k_\w+)(<[^>]*>)?", name)
    key = (m.group(1) + (m.group(2) or "")) if m else name[:40]
    a = agg[key]; a[0] += 1; a[1] += ns; a[2] = max(a[2], ns)
print(f"{len(rows)} launches, total {tot/1e6:.2f} ms (serialised, cold-cache)")
print(f"{'kernel':48s} {'launches':>8s} {'total ms':>10s} {'share':>7s} {'avg us':>9s} {'max us':>9s}")
for key, (n, ns, mx) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{key:48s} {n:8d} {ns/1e6:10.3f} {100*ns/tot:6.1f}% {ns/n/1e3:9.1f} {mx/1e3:9.1f}")
