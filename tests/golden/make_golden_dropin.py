"""Generates tests/golden/golden_dropin.npz by running the reference's own per-frame functions of
depth_map.py (apply_stereo_rectification, create_depth_map; imported from /root/reference) on a seeded
synthetic BGR pair with the synthetic calibration of tests/golden_cases.synth_calibration.
Runs in the build container only (needs /root/reference and cv2):
    python tests/golden/make_golden_dropin.py
"""
import contextlib
import io
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
sys.path.insert(0, ROOT)
from stereovision_b200.synth import make_pair, to_bgr  # noqa: E402
from tests.golden_cases import DROPIN_CASE, synth_calibration  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    sys.path.insert(0, "/root/reference")
    with contextlib.redirect_stdout(io.StringIO()):
        import depth_map as ref
    c = DROPIN_CASE
    l, r, _ = make_pair(c["H"], c["W"], 64, seed=c["seed"])
    lb, rb = to_bgr(l), to_bgr(r)
    calib = synth_calibration(c["W"], c["H"])
    with contextlib.redirect_stdout(io.StringIO()):
        lr, rr = ref.apply_stereo_rectification(lb, rb, calib)
        depth, disp, _ = ref.create_depth_map(lr, rr, None)
        depth_raw, disp_raw, _ = ref.create_depth_map(lb, rb, None, min_depth=0.2, max_depth=40.0)
    out = dict(left_map1=calib["left_map1"], left_map2=calib["left_map2"], right_map1=calib["right_map1"],
               right_map2=calib["right_map2"], left_rect=lr, right_rect=rr, depth=depth, disparity=disp,
               depth_raw=depth_raw, disparity_raw=disp_raw,
               params=np.array([ref.MIN_DISP, ref.NUM_DISP, ref.WINDOW_SIZE]))
    path = os.path.join(HERE, "golden_dropin.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes", depth.dtype, disp.dtype, float((depth > 0).mean()))


if __name__ == "__main__":
    main()
