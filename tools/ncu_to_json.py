#!/usr/bin/env python
"""Key metrics of one kernel of an .ncu-rep as JSON (read by bench.py for roofline.traffic).
usage: ncu_to_json.py rep.ncu-rep out.json frames H W sf [kernel-substring]"""
import csv
import io
import json
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]
wl = [int(a) for a in sys.argv[3:7]]
pat = sys.argv[7] if len(sys.argv) > 7 else ""
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
row = [r for r in rows[2:] if pat in r[hdr.index("Kernel Name")]][-1]


def get(name):
    v, u = float(row[hdr.index(name)].replace(",", "")), units[hdr.index(name)]
    return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "us": 1, "ms": 1e3, "ns": 1e-3}.get(u, 1)


j = {"workload": wl, "kernel": row[hdr.index("Kernel Name")][:120], "source": rep.split("/")[-1],
     "dram_bytes_read": int(get("dram__bytes_read.sum")), "dram_bytes_write": int(get("dram__bytes_write.sum")),
     "duration_us_under_ncu": get("gpu__time_duration.sum"),
     "dram_throughput_pct": get("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
     "issue_active_pct": get("smsp__issue_active.avg.pct_of_peak_sustained_active"),
     "warps_active_pct": get("sm__warps_active.avg.pct_of_peak_sustained_active"),
     "registers_per_thread": int(get("launch__registers_per_thread"))}
json.dump(j, open(out, "w"), indent=1)
print(json.dumps(j))
