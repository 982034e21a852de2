#!/usr/bin/env python
"""Dynamic SASS mix of one kernel from `ncu --page source --csv` output:
python tools/ncu_sass_mix.py source.csv <units>   (units = loop trips to normalise by, e.g. warp-columns)"""
import csv, re, sys
from collections import defaultdict
rows = list(csv.reader(open(sys.argv[1])))
ends = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
if len(ends) > 1:
    rows = rows[:ends[1]]            # the export repeats the listing; keep the first copy
units = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
hdr = rows[1]
ia, isrc, iex, ismp = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_")]
tot = 0; by = defaultdict(float); smp = defaultdict(float); st = defaultdict(float)
for r in rows[2:]:
    if len(r) <= iex or not r[iex] or not r[iex].replace(".", "").isdigit():
        continue
    n = float(r[iex]); tot += n
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_x]+)*)", r[isrc])
    op = m.group(2) if m else "?"
    key = op.split(".")[0] + ("." + ".".join(p for p in op.split(".")[1:] if p in ("U16x2", "S16x2", "16x2", "128", "64", "MIN", "MOV", "IADD", "SHL", "WIDE")) if "." in op else "")
    by[key] += n; smp[key] += float(r[ismp] or 0)
    for i in stall_cols:
        if r[i]:
            st[hdr[i]] += float(r[i])
print(f"total warp-instr {tot:.4g}  per unit {tot / units:.1f}")
for k, v in sorted(by.items(), key=lambda kv: -kv[1])[:40]:
    print(f"  {k:28s} {v / units:8.2f} /unit   samples {smp[k]:8.0f}")
print({k: int(v) for k, v in sorted(st.items(), key=lambda kv: -kv[1])})
