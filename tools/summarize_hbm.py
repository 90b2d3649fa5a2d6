"""Summarise an ncu `--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv` launch list:
per kernel, DRAM bytes moved over its launch time = achieved HBM GB/s, against the measured copy bandwidth in
MEASURED_PEAKS.json (B200_PROFILING.md recipe).  Usage: python tools/summarize_hbm.py launches.csv [peak_gbs]"""
import csv, json, re, sys
from collections import defaultdict
from pathlib import Path

peak = float(sys.argv[2]) if len(sys.argv) > 2 else None
if peak is None:
    p = Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json"
    peak = float(json.loads(p.read_text())["hbm_gbs"]) if p.exists() else 6650.0
rows = defaultdict(dict)   # launch id -> {name, metric: value}
with open(sys.argv[1]) as fh:
    for line in fh:
        if line.startswith('"ID"'):
            header = next(csv.reader([line]))
            break
    idx = {h: i for i, h in enumerate(header)}
    for r in csv.reader(fh):
        if len(r) < len(header):
            continue
        d = rows[r[idx["ID"]]]
        d["name"] = r[idx["Kernel Name"]]
        unit, val = r[idx["Metric Unit"]], float(r[idx["Metric Value"]].replace(",", ""))
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1.0, "us": 1e3, "ms": 1e6, "nsecond": 1.0, "usecond": 1e3, "msecond": 1e6}.get(unit, 1.0)
        d[r[idx["Metric Name"]]] = val * scale
agg = defaultdict(lambda: [0, 0.0, 0.0, 0.0])
for d in rows.values():
    m = re.search(r"(k_\w+)(<[^>]*>)?", d["name"])
    key = (m.group(1) + (m.group(2) or "")) if m else d["name"][:40]
    a = agg[key]
    a[0] += 1
    a[1] += d.get("gpu__time_duration.sum", 0.0)
    a[2] += d.get("dram__bytes_read.sum", 0.0)
    a[3] += d.get("dram__bytes_write.sum", 0.0)
print(f"HBM peak used: {peak:.1f} GB/s; per-launch times are ncu's (serialised, cold cache)")
print(f"{'kernel':44s} {'launches':>8s} {'total ms':>9s} {'read MB':>10s} {'write MB':>10s} {'GB/s':>8s} {'of peak':>8s}")
for key, (n, ns, rd, wr) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    gbs = (rd + wr) / ns if ns > 0 else 0.0
    print(f"{key:44s} {n:8d} {ns/1e6:9.3f} {rd/1e6:10.1f} {wr/1e6:10.1f} {gbs:8.1f} {100*gbs/peak:7.1f}%")
