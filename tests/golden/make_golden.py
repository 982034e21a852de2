"""Generates tests/golden/*.npz with the reference's own arithmetic: live OpenCV
(opencv-python-headless 4.13.0.92), the third-party dependency the reference calls on its
per-frame path (depth_map.py:815-826, 871-912; BASELINE.json's StereoBM / reprojectImageTo3D).
Inputs are re-created in the tests from seeds (stereovision_b200.synth); only parameters and
cv2's outputs are stored.  Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import cv2
import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from stereovision_b200.synth import make_pair, to_bgr  # noqa: E402
from tests.golden_cases import (SGBM_CASES, SGBM_COLOR_CASES, BM_CASES, RESIZE_CASES, color_pair, synth_calibration,  # noqa: E402
                                resize_case_input)

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    out = {"cv2_version": np.array(cv2.__version__)}
    for i, c in enumerate(SGBM_CASES):
        l, r, _ = make_pair(c["H"], c["W"], max(c["prm"]["numDisparities"], 16), seed=c["seed"])
        out[f"sgbm_{i}"] = cv2.StereoSGBM_create(**c["prm"]).compute(l, r)
    for i, c in enumerate(SGBM_COLOR_CASES):
        l, r = color_pair(c)
        out[f"sgbm_color_{i}"] = cv2.StereoSGBM_create(**c["prm"]).compute(l, r)
    for i, c in enumerate(BM_CASES):
        l, r, _ = make_pair(c["H"], c["W"], max(c["numDisparities"], 16), seed=c["seed"], noise=c.get("noise", 6))
        bm = cv2.StereoBM_create(c["numDisparities"], c["blockSize"])
        for k, v in c.get("set", {}).items():
            getattr(bm, "set" + k)(v)
        out[f"bm_{i}"] = bm.compute(l, r)
    # cv2.validateDisparity on a BM disparity with seeded int16 / int32 cost maps
    c = BM_CASES[0]
    l, r, _ = make_pair(c["H"], c["W"], 16, seed=c["seed"], noise=25)
    bm = cv2.StereoBM_create(c["numDisparities"], c["blockSize"]); bm.setUniquenessRatio(0); bm.setTextureThreshold(0)
    d = bm.compute(l, r)
    rng = np.random.default_rng(77)
    out["validate_in"] = d
    out["validate_cost16"] = rng.integers(0, 40, d.shape).astype(np.int16)      # many ties: lowest x must win
    out["validate_cost32"] = rng.integers(-100000, 100000, d.shape).astype(np.int32)
    for k, (cost, d12) in enumerate(((out["validate_cost16"], 1), (out["validate_cost32"], 0), (out["validate_cost16"], 4))):
        v = d.copy()
        cv2.validateDisparity(v, cost, 0, c["numDisparities"], d12)
        out[f"validate_out_{k}"] = v
    # remap + gray with maps from the reference's own recipe (depth_map.py:625-638)
    H, W = 120, 160
    calib = synth_calibration(W, H)
    l, r, _ = make_pair(H, W, 16, seed=7)
    bgr = to_bgr(l)
    out["remap_map1"], out["remap_map2"] = calib["left_map1"], calib["left_map2"]
    out["remap_bgr"] = cv2.remap(bgr, calib["left_map1"], calib["left_map2"], cv2.INTER_LINEAR)
    out["remap_gray"] = cv2.remap(l, calib["left_map1"], calib["left_map2"], cv2.INTER_LINEAR)
    out["gray"] = cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY)
    out["Q"] = calib["Q"]
    # median / speckle / reproject on a cv2 disparity
    prm = SGBM_CASES[0]["prm"]
    l0, r0, _ = make_pair(SGBM_CASES[0]["H"], SGBM_CASES[0]["W"], 16, seed=SGBM_CASES[0]["seed"])
    d16 = cv2.StereoSGBM_create(**{**prm, "speckleWindowSize": 0}).compute(l0, r0)
    out["med3_in"] = d16
    out["med3"] = cv2.medianBlur(d16, 3)
    f = d16.astype(np.float32) / 16.0
    out["med5"] = cv2.medianBlur(f, 5)
    sp = d16.copy()
    cv2.filterSpeckles(sp, (prm["minDisparity"] - 1) * 16, 40, 16)
    out["speckle"] = sp
    Hh, Ww = d16.shape
    Q = synth_calibration(Ww, Hh)["Q"]
    out["Q2"] = Q
    out["reproj_s16"] = cv2.reprojectImageTo3D(d16, Q)
    out["reproj_f32"] = cv2.reprojectImageTo3D(f, Q, handleMissingValues=True)
    # initUndistortRectifyMap (depth_map.py:633-638): the reference's own recipe on the synthetic calibration,
    # plus the 8- and 12-coefficient distortion models
    Wm, Hm = 200, 140
    cal = synth_calibration(Wm, Hm)
    R1, R2, P1, P2, _, _, _ = cv2.stereoRectify(cal["mtx_left"], cal["dist_left"], cal["mtx_right"], cal["dist_right"],
                                                (Wm, Hm), cal["R"], cal["T"], alpha=0)
    d8 = np.array([-0.3, 0.1, 2e-3, -1e-3, 0.02, 0.01, -0.02, 0.005])
    d12 = np.concatenate([d8, [1e-3, -2e-3, 5e-4, 1e-4]])
    out["maps_K"], out["maps_R1"], out["maps_P1"], out["maps_R2"], out["maps_P2"] = cal["mtx_left"], R1, P1, R2, P2
    for tag, (K, D, R, P) in {"l5": (cal["mtx_left"], cal["dist_left"], R1, P1),
                              "r5": (cal["mtx_right"], cal["dist_right"], R2, P2),
                              "l8": (cal["mtx_left"], d8, R1, P1), "l12": (cal["mtx_left"], d12, R1, P1),
                              "l0": (cal["mtx_left"], None, None, cal["mtx_left"])}.items():
        m1, m2 = cv2.initUndistortRectifyMap(K, D, R, P, (Wm, Hm), cv2.CV_16SC2)
        out[f"maps_{tag}_m1"], out[f"maps_{tag}_m2"] = m1, m2
        out[f"maps_{tag}_dist"] = np.zeros(0) if D is None else np.asarray(D, np.float64)
    # cv2.resize(INTER_LINEAR): the reference's per-frame downscale and its float32 mono-map resize
    for i, c in enumerate(RESIZE_CASES):
        out[f"resize_{i}"] = cv2.resize(resize_case_input(c), c["dsize"], interpolation=cv2.INTER_LINEAR)
    np.savez_compressed(os.path.join(HERE, "golden.npz"), **out)
    print("wrote", os.path.join(HERE, "golden.npz"), os.path.getsize(os.path.join(HERE, "golden.npz")), "bytes")


if __name__ == "__main__":
    main()
