"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.

    python profiles/make_launch_summary.py gpurun_out/r03b_launches.csv "<command that was profiled>" > profiles/r03b_launches_summary.txt
"""
import csv
import re
import sys
from collections import defaultdict

path, what = sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else ""
rows = []
with open(path, newline="") as f:
    lines = [ln for ln in f if ln.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    ms = v * {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "nsecond": 1e-6, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3}[unit]
    name = re.sub(r"\(.*", "", r["Kernel Name"])[:60]
    rows.append((name, ms))
tot = sum(ms for _, ms in rows)
agg = defaultdict(lambda: [0, 0.0])
for n, ms in rows:
    agg[n][0] += 1
    agg[n][1] += ms
print(f"# ncu --metrics gpu__time_duration.sum --clock-control none -c 400: {what}; {len(rows)} launches, {tot:.1f} ms of device time")
for n, (c, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{n:60s} {c:6d} launches {ms:12.3f} ms {100 * ms / tot:6.2f} %")
own = sum(ms for n, (c, ms) in agg.items()
          if "dg::" in n or "dgw::" in n or "dgn::" in n or "oz::oz_" in n  # oz:: = dgw::oz (csrc/ozaki_gemm.cuh; ncu prints the inner namespace)
          or n.startswith("<unnamed>::") or n.startswith("void <unnamed>::"))  # (anonymous namespaces of csrc/)
print(f"# kernels of this repository: {100 * own / tot:.2f} % of the device time")
