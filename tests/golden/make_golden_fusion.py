"""Generates tests/golden/golden_fusion.npz by IMPORTING the reference's own functions from
/root/reference/fused_depth_map.py (create_depth_map_stereo_scaled, calibrate_midas_to_stereo,
normalize_to_stereo_range, fuse_depth_maps, detect_camera_occlusion) and running them on the
seeded synthetic inputs of tests/golden_cases.FUSION_CASES.  Only outputs are stored; the tests
re-create the inputs from the seeds.  Needs /root/reference and cv2, so it runs in the build
container only:   python tests/golden/make_golden_fusion.py
"""
import contextlib
import io
import os
import sys

import cv2
import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
sys.path.insert(0, ROOT)
from tests.golden_cases import FUSION_CASES, fusion_case_inputs as case_inputs  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    sys.path.insert(0, "/root/reference")
    with contextlib.redirect_stdout(io.StringIO()):
        import fused_depth_map as ref
    out = {"cv2_version": np.array(cv2.__version__)}
    defaults = dict(ref.fusion_params)
    for i, c in enumerate(FUSION_CASES):
        lb, rb, midas, flow = case_inputs(c)
        with contextlib.redirect_stdout(io.StringIO()):
            norm, disp, _, conf = ref.create_depth_map_stereo_scaled(lb, rb, c["min_disp"], c["num_disp"], c["window"])
            cal = ref.calibrate_midas_to_stereo(midas, disp, conf)
            cal_few = ref.calibrate_midas_to_stereo(midas, disp, np.zeros_like(conf))
            # a mono map of another shape: the reference resizes it first (fused_depth_map.py:1214-1217)
            cal_rs = ref.calibrate_midas_to_stereo(np.ascontiguousarray(midas[::2, ::3]), disp, conf)
            flow_n = ref.normalize_to_stereo_range(flow, disp)
            ref.fusion_params.update(defaults)
            ref.fusion_params.update(c["params"])
            fused, _ = ref.fuse_depth_maps(norm, conf, cal, None, flow_n, c["moving"])
            fused_m, _ = ref.fuse_depth_maps(None, None, cal, None, flow_n, c["moving"], use_stereo=False)
            fused_f, _ = ref.fuse_depth_maps(None, None, None, None, flow_n, True, use_stereo=False, use_midas=False)
            occ = ref.detect_camera_occlusion(lb, rb)
        out[f"norm_{i}"] = norm.astype(np.uint8)
        out[f"disp_{i}"] = disp
        out[f"conf_{i}"] = conf.astype(np.uint8)
        out[f"cal_{i}"] = cal
        out[f"calfew_{i}"] = cal_few
        out[f"calrs_{i}"] = cal_rs
        out[f"flown_{i}"] = flow_n
        out[f"fused_{i}"] = fused
        out[f"fusedm_{i}"] = fused_m
        out[f"fusedf_{i}"] = fused_f
        out[f"occ_{i}"] = np.array([{"none": 0, "left": 1, "right": 2, "both": 3}[occ[0]], occ[1], occ[2]], np.float64)
    ref.fusion_params.update(defaults)
    # the two OpenCV filters on their own (pins oracle.fusion.gaussian_blur / bilateral_filter)
    lb, rb, midas, flow = case_inputs(FUSION_CASES[0])
    out["gauss15"] = cv2.GaussianBlur(flow, (15, 15), 0)
    out["bilat9"] = cv2.bilateralFilter(flow, 9, 75, 75)
    path = os.path.join(HERE, "golden_fusion.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
