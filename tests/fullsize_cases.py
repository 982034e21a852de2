"""Case tables for the full-size parity pins (tests/golden/make_golden_fullsize.py runs cv2 on them;
the tests and bench.py only read the committed hashes).  BASELINE.json configs:
config 1 = 640x480 D=64, config 3 = 1280x720 D=128 MODE_SGBM, config 4 = 1920x1080 D=256 MODE_HH;
"script" = the parameters depth_map.py itself uses (depth_map.py:894-906: D=320, bs=7, 3-channel P1/P2, 3WAY)."""
import hashlib
import json
import os

import numpy as np


def _p(D, bs, mode, p1=None, p2=None):
    return dict(minDisparity=0, numDisparities=D, blockSize=bs, P1=8 * bs * bs if p1 is None else p1,
                P2=32 * bs * bs if p2 is None else p2, disp12MaxDiff=1, preFilterCap=63, uniquenessRatio=10,
                speckleWindowSize=100, speckleRange=32, mode=mode)


FULLSIZE_CASES = {
    "c4_1080p_d256_hh": dict(H=1080, W=1920, seeds=[1, 2], prm=_p(256, 5, 1)),
    "c3_720p_d128_sgbm": dict(H=720, W=1280, seeds=[1, 2], prm=_p(128, 5, 0)),
    "script_480p_d320_3way": dict(H=480, W=640, seeds=[1, 2], prm=_p(320, 7, 2, p1=8 * 3 * 49, p2=32 * 3 * 49)),
    "c1_480p_d64_3way": dict(H=480, W=640, seeds=[1], prm=_p(64, 5, 2)),
    "c1_480p_d64_sgbm": dict(H=480, W=640, seeds=[1], prm=_p(64, 5, 0)),
    "hh4_720p_d128": dict(H=720, W=1280, seeds=[1], prm=_p(128, 5, 3)),
}

# bench.py's workloads (same table as bench.WORKLOADS; tests/test_bench_contract.py asserts they agree)
BENCH_WORKLOADS = {
    "sgbm_1080p_d256_hh": dict(H=1080, W=1920, D=256, mode=1),
    "sgbm_720p_d128_sgbm": dict(H=720, W=1280, D=128, mode=0),
    "sgbm_480p_d64_sgbm": dict(H=480, W=640, D=64, mode=0),
    "sgbm_480p_d320_3way": dict(H=480, W=640, D=320, mode=2),
}
BENCH_RANKS = 8       # ranks whose seeds (1000*rank + i) are pinned for the headline workload
BENCH_SCENES = 4      # distinct scenes per rank (bench.make_inputs tiles them over the batch)


def bench_params(wl):
    return _p(wl["D"], 5, wl["mode"])


_HERE = os.path.dirname(os.path.abspath(__file__))


def load_hashes():
    with open(os.path.join(_HERE, "golden", "fullsize_sha256.json")) as f:
        return json.load(f)


def sha256_of(d):
    return hashlib.sha256(np.ascontiguousarray(d).tobytes()).hexdigest()
