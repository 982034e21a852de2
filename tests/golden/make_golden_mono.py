"""Generates tests/golden/golden_mono.npz by running the reference's own estimator classes, imported from
/root/reference/fused_depth_map.py, with only the parts this repo does not cover replaced:
  * DepthEstimatorMidas.__call__ runs unmodified on an instance whose `model` returns the seeded synthetic
    prediction of tests/golden_cases.mono_case_inputs (no network weights are available offline) -- so the
    bicubic upsample (torch), the normalisation and the contrast confidence (cv2) are the reference's;
  * OpticalFlowDepthEstimator.__call__ runs unmodified on seeded panning frames; cv2.calcOpticalFlowFarneback
    and cv2.findHomography are wrapped only to RECORD their results (the dense flow and the homography are the
    inputs of this repo's flow_residual_depth), everything after them is the reference's.
Runs in the build container only (needs /root/reference, cv2, torch):  python tests/golden/make_golden_mono.py
"""
import contextlib
import io
import os
import sys

import cv2
import numpy as np
import torch

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
sys.path.insert(0, ROOT)
from tests.golden_cases import MONO_CASES, mono_case_inputs, mono_flow_frames  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    sys.path.insert(0, "/root/reference")
    with contextlib.redirect_stdout(io.StringIO()):
        import fused_depth_map as ref
    out = {"cv2_version": np.array(cv2.__version__), "torch_version": np.array(torch.__version__)}
    for i, c in enumerate(MONO_CASES):
        frame, pred = mono_case_inputs(c)
        est = object.__new__(ref.DepthEstimatorMidas)
        est.device = torch.device("cpu")
        est.transform = lambda img: torch.zeros(1, 3, 8, 8)
        est.model = lambda batch, p=pred: torch.from_numpy(p)[None]
        with contextlib.redirect_stdout(io.StringIO()):
            depth, conf = est(frame)
        assert depth is not None and depth.shape == frame.shape[:2]
        out[f"midas_depth_{i}"] = depth
        out[f"midas_conf_{i}"] = conf
        if c.get("flat"):
            continue
        # the flow estimator over a few frames: the first call primes prev_gray.  Before the last frame the
        # motion timeout is made to expire: the valid-homography branch sets camera_moving = True before it
        # picks alpha (fused_depth_map.py:1381-1396), so the average still uses 0.9 -- recorded as such.
        rec = {}
        real_flow, real_homo = cv2.calcOpticalFlowFarneback, cv2.findHomography

        def flow_spy(*a, **k):
            rec["flow"] = real_flow(*a, **k)
            return rec["flow"]

        def homo_spy(*a, **k):
            rec["H"], rec["mask"] = real_homo(*a, **k)
            return rec["H"], rec["mask"]
        cv2.calcOpticalFlowFarneback, cv2.findHomography = flow_spy, homo_spy
        try:
            with contextlib.redirect_stdout(io.StringIO()):
                fe = ref.OpticalFlowDepthEstimator()
                frames = mono_flow_frames(c)
                assert fe(frames[0]) is None
                for k, f in enumerate(frames[1:]):
                    if k == 2:
                        fe.last_move_time -= 10.0
                    stable_in = None if fe.stable_depth is None else fe.stable_depth.copy()
                    moving_in = True
                    sm = fe(f)
                    assert sm is not None, "homography rejected"
                    out[f"flow_{i}_{k}"] = rec["flow"]
                    out[f"homography_{i}_{k}"] = np.asarray(rec["H"], np.float64)
                    out[f"flow_moving_{i}_{k}"] = np.array(moving_in)
                    out[f"flow_smoothed_{i}_{k}"] = sm
                    out[f"flow_stable_{i}_{k}"] = fe.stable_depth.copy()
                    assert stable_in is None or stable_in.shape == sm.shape
        finally:
            cv2.calcOpticalFlowFarneback, cv2.findHomography = real_flow, real_homo
    path = os.path.join(HERE, "golden_mono.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
