"""Generates tests/golden/golden_calib.npz by running the reference's OWN calibration loaders, unmodified,
on a synthetic 7-key pickle written where they look for it (output/stereo_calibration_data.pkl relative to the
working directory, depth_map.py:22):

    depth_map.load_stereo_calibration()                                depth_map.py:591-668
    fused_depth_map.load_stereo_calibration_with_scaling(s), s = 1.0, 0.5, 0.33   fused_depth_map.py:307-442

Stored: both CV_16SC2 map pairs, Q, ROIs, sizes, focal length and baseline of every call.  Runs in the build
container only (needs /root/reference and cv2):  python tests/golden/make_golden_calib.py
"""
import contextlib
import io
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
sys.path.insert(0, ROOT)
from stereovision_b200.synth import write_calibration_pkl  # noqa: E402
from tests.golden_cases import CALIB_CASE  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def pack(out, tag, d, size_key):
    for k in ("left_map1", "left_map2", "right_map1", "right_map2", "Q"):
        out[f"{tag}_{k}"] = np.asarray(d[k])
    out[f"{tag}_roi"] = np.array([d["roi1"], d["roi2"]], np.int64)
    out[f"{tag}_size"] = np.array(d[size_key], np.int64)
    out[f"{tag}_baseline"] = np.float64(d["baseline"])
    if "focal_length" in d:
        out[f"{tag}_focal_length"] = np.float64(d["focal_length"])
        out[f"{tag}_mtx_left"] = np.asarray(d["mtx_left"])


def main():
    sys.path.insert(0, "/root/reference")
    cwd = os.getcwd()
    out = {}
    with tempfile.TemporaryDirectory() as tmp:
        os.chdir(tmp)
        try:
            write_calibration_pkl(os.path.join("output", "stereo_calibration_data.pkl"), CALIB_CASE["W"], CALIB_CASE["H"])
            with contextlib.redirect_stdout(io.StringIO()):
                import depth_map as ref
                import fused_depth_map as fref
                pack(out, "plain", ref.load_stereo_calibration(), "img_size")
                for s in CALIB_CASE["scales"]:
                    pack(out, f"scale{int(round(s * 100))}", fref.load_stereo_calibration_with_scaling(s), "img_size_proc")
        finally:
            os.chdir(cwd)
    path = os.path.join(HERE, "golden_calib.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes", sorted(out)[:6])


if __name__ == "__main__":
    main()
