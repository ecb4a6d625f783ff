"""profiles/r02_traffic.json from an ncu launch list (the CSV of tools/prof_r02.sh): DRAM read+write bytes per launch,
averaged per kernel, stamped with the hash of the kernel sources it was measured on (bench.source_stamp), so that
bench.py drops `roofline.traffic` when the kernels have changed since.
    python tools/traffic_from_launches.py gpurun_out/r02_launches.csv profiles/r02_traffic.json"""
import csv
import json
import os
import re
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr, data = rows[hi], rows[hi + 1:]
ki, vi, mi, ii = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Name"), hdr.index("ID")
per = {}
for r in data:
    if len(r) <= vi:
        continue
    d = per.setdefault(int(r[ii]), {"name": r[ki]})
    d[r[mi]] = float(r[vi].replace(",", ""))
agg = {}
for d in per.values():
    m = re.search(r"(\w+_kernel)", d["name"])
    if not m:
        continue
    a = agg.setdefault(m.group(1), {"bytes": 0.0, "ns": 0.0, "n": 0})
    a["bytes"] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
    a["ns"] += d.get("gpu__time_duration.sum", 0.0)
    a["n"] += 1
out = {"source_stamp": bench.source_stamp(), "from": os.path.basename(sys.argv[1]),
       "per_launch_bytes": {k: v["bytes"] / v["n"] for k, v in agg.items()},
       "per_launch_us": {k: v["ns"] / v["n"] / 1e3 for k, v in agg.items()},
       "launches": {k: v["n"] for k, v in agg.items()}}
json.dump(out, open(sys.argv[2], "w"), indent=1)
print(json.dumps(out))
