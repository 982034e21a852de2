#!/usr/bin/env python
"""A/B helper for kernel experiments: run bench.py once per library variant (build/libsvb_<tag>.so, made by
tools/build_variant.sh; 'base' = the in-tree library) and print one line per variant:
pairs/s, parity, per-stage ms.   python tools/ab_bench.py base r16 r12 [-- extra bench args]"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
args = sys.argv[1:]
extra = []
if "--" in args:
    i = args.index("--")
    args, extra = args[:i], args[i + 1:]
for tag in args:
    env = dict(os.environ)
    if tag != "base":
        env["SVB_LIBRARY"] = os.path.join(ROOT, "build", f"libsvb_{tag}.so")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "5", "--warmup", "2",
                        "--no-cpu-baseline"] + extra, env=env, capture_output=True, text=True)
    try:
        d = json.loads(p.stdout.strip().splitlines()[-1])
        st = {k: round(v, 2) for k, v in d["roofline"]["stages_ms_per_step"].items()}
        pg = (d.get("e2e_pageable") or {}).get("value")
        print(f"{tag:10s} {d['value']:8.1f} pairs/s  e2e {d['e2e']['value']:8.1f}  pageable {pg if pg is None else round(pg, 1)}  "
              f"parity {d['parity']['sha256_match']}  {st}", flush=True)
    except Exception as e:
        print(tag, "FAILED", e, p.stderr[-1500:], flush=True)
