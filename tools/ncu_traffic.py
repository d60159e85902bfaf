#!/usr/bin/env python3
"""DRAM traffic per launch of the two region kernels from an `ncu --set full` report.

usage: ncu_traffic.py report.ncu-rep workload frames_per_launch [out.json]
Writes profiles/traffic.json, which bench.py reads for roofline.traffic."""
import csv
import io
import json
import os
import subprocess
import sys

rep, workload, frames = sys.argv[1], sys.argv[2], int(sys.argv[3])
out_path = sys.argv[4] if len(sys.argv) > 4 else os.path.join(
    os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "traffic.json")
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
h, u = rows[0], rows[1]
mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
res = {"workload": workload, "frames_per_launch": frames,
       "source": "ncu --set full --clock-control none, %s" % os.path.basename(rep)}
for r in rows[2:]:
    d = dict(zip(h, r))
    name = d["Kernel Name"].split("(")[0].split("::")[-1]
    tot = 0.0
    for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        tot += float(d[k].replace(",", "")) * mult[u[h.index(k)]]
    res[name] = tot
    res[name + "_time_us"] = d.get("gpu__time_duration.sum")
json.dump(res, open(out_path, "w"), indent=1)
print(json.dumps(res, indent=1))
