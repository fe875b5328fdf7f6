"""Print kernel name + duration from an `ncu --metrics gpu__time_duration.sum --csv` log.

    python tools/launch_list.py gpurun_out/launches.csv [last_n]
"""
import csv
import sys

path = sys.argv[1]
last = int(sys.argv[2]) if len(sys.argv) > 2 else 0
with open(path) as f:
    lines = [l for l in f if l.startswith('"')]
rows = [(d["Kernel Name"], float(d["Metric Value"].replace(",", "")), d["Metric Unit"])
        for d in csv.DictReader(lines) if d.get("Metric Name") == "gpu__time_duration.sum"]
if last:
    rows = rows[-last:]
total = 0.0
for name, v, unit in rows:
    us = v / 1000.0 if unit == "ns" else v
    total += us
    print(f"{us:9.1f} us  {name[:110]}")
print(f"{total:9.1f} us  total of {len(rows)} launches")
