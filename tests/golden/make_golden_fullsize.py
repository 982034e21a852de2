"""Full-size parity pins: SHA-256 of cv2.StereoSGBM.compute (opencv-python-headless 4.13.0.92, the
arithmetic the reference calls at depth_map.py:894-909 / fused_depth_map.py:988-1004) on seeded
synthetic pairs at the BASELINE.json sizes, written to tests/golden/fullsize_sha256.json.

A 1080p int16 disparity is 4 MB, so only its hash (plus a few summary numbers that make a mismatch
debuggable) is committed; the inputs are re-created from the seed.  Two groups:

  * "tests": one or two seeds per graded config (config 3, config 4, the script's own 640x480 / D=320 /
    bs=7 / 3WAY, config 1) -- compared by tests/test_fullsize_parity_gpu.py and (oracle leg) by
    tests/test_fullsize_oracle.py;
  * "bench": the seeds bench.py feeds each rank (seed0 = 1000*rank, four distinct scenes per rank), so the
    bench line can prove its own output outside the timed region.

Run from the repo root (about 3 minutes of cv2):  python tests/golden/make_golden_fullsize.py
"""
import hashlib
import json
import os
import sys

import cv2
import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from stereovision_b200.synth import make_pair  # noqa: E402
from tests.fullsize_cases import FULLSIZE_CASES, BENCH_WORKLOADS, bench_params, BENCH_RANKS, BENCH_SCENES  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def describe(d):
    d = np.ascontiguousarray(d)
    return {"sha256": hashlib.sha256(d.tobytes()).hexdigest(), "valid": int((d >= 0).sum()),
            "sum": int(d.astype(np.int64).sum())}


def main():
    cv2.setNumThreads(1)
    out = {"cv2_version": cv2.__version__, "tests": {}, "bench": {}}
    for name, c in FULLSIZE_CASES.items():
        for seed in c["seeds"]:
            l, r, _ = make_pair(c["H"], c["W"], c["prm"]["numDisparities"], seed=seed)
            d = cv2.StereoSGBM_create(**c["prm"]).compute(l, r)
            out["tests"][f"{name}/seed{seed}"] = describe(d)
            print(name, seed, out["tests"][f"{name}/seed{seed}"]["sha256"][:16], flush=True)
    for name, wl in BENCH_WORKLOADS.items():
        prm = bench_params(wl)
        for rank in range(BENCH_RANKS if name == "sgbm_1080p_d256_hh" else 1):
            for i in range(BENCH_SCENES):
                seed = 1000 * rank + i
                l, r, _ = make_pair(wl["H"], wl["W"], wl["D"], seed=seed)
                d = cv2.StereoSGBM_create(**prm).compute(l, r)
                out["bench"][f"{name}/seed{seed}"] = describe(d)
                print("bench", name, seed, out["bench"][f"{name}/seed{seed}"]["sha256"][:16], flush=True)
    with open(os.path.join(HERE, "fullsize_sha256.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
