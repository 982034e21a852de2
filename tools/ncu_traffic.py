#!/usr/bin/env python
"""Per-kernel DRAM traffic and duration of an `ncu --set full` report as JSON (bench.py's roofline.traffic reads
the committed file instead of constants):
    python tools/ncu_traffic.py gpurun_out/r02_v1_b32.ncu-rep sgbm_1080p_d256_hh 32 > profiles/r02_ncu_traffic.json"""
import csv
import io
import json
import re
import subprocess
import sys

rep, workload, pairs = sys.argv[1], sys.argv[2], int(sys.argv[3])
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]


def val(r, name):
    i = hdr.index(name)
    v = float(r[i].replace(",", ""))
    u = units[i].lower()
    scale = {"gbyte": 1e9, "mbyte": 1e6, "kbyte": 1e3, "byte": 1.0, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1.0,
             "msecond": 1e-3, "usecond": 1e-6, "nsecond": 1e-9, "second": 1.0}.get(u, 1.0)
    return v * scale


kernels = {}
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")]
    m = re.match(r"(?:void )?(k_\w+)(<[^>]*>)?", name)
    key = m.group(1) if m else name
    if key == "k_sweep":
        key = "k_sweep<last+wta>" if re.search(r",\s*(1|true)\s*>", m.group(2) or "") else "k_sweep<first>"
    kernels[key] = {"dram_bytes_read": val(r, "dram__bytes_read.sum"), "dram_bytes_write": val(r, "dram__bytes_write.sum"),
                    "seconds_under_ncu": val(r, "gpu__time_duration.sum"), "kernel": name}
    kernels[key]["dram_bytes_per_pair"] = (kernels[key]["dram_bytes_read"] + kernels[key]["dram_bytes_write"]) / pairs
json.dump({"source": rep.split("/")[-1], "workload": workload, "pairs_per_launch": pairs, "kernels": kernels}, sys.stdout, indent=1)
print()
