#!/usr/bin/env python
"""Print the headline fields of bench.py JSON lines:  python tools/print_bench.py gpurun_out/b.json [more.json ...]"""
import json
import sys

for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable:", e)
        continue
    r = d.get("roofline") or {}
    print(f, "| value", round(d.get("value", 0), 1), d.get("unit"), "| e2e", round((d.get("e2e") or {}).get("value", 0), 1),
          "| pageable", round((d.get("e2e_pageable") or {}).get("value", 0), 1),
          "| parity", (d.get("parity") or {}).get("sha256_match"), "| frac", round(r.get("frac", 0), 4),
          "| fullpath", {k: round(v, 1) for k, v in (d.get("fullpath") or {}).items() if isinstance(v, float)},
          "| stages", {k: round(v, 2) for k, v in (r.get("stages_ms_per_step") or {}).items()})
