#!/usr/bin/env python
"""One line per frame configuration of a tools/bench_frame.py JSON (stdin or file): call-by-call ms, svb.frame ms, cv2 ms."""
import json
import sys

txt = open(sys.argv[1]).read() if len(sys.argv) > 1 else sys.stdin.read()
d = json.loads(txt.strip().splitlines()[-1])
print({k: (round(v["svb_ms_per_frame"], 2), None if v.get("svb_frame_api_ms_per_frame") is None else round(v["svb_frame_api_ms_per_frame"], 2),
           round(v["cv2_ms_per_frame"], 1)) for k, v in d["frames"].items()})
