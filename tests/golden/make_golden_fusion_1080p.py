"""BASELINE.json config 5 at its own size: generates tests/golden/golden_fusion_1080p.npz by running the reference's
OWN functions (imported from /root/reference/fused_depth_map.py) at 1920 x 1080:

    create_depth_map_stereo_scaled (D = 256, window 5) -> calibrate_midas_to_stereo -> normalize_to_stereo_range
    -> fuse_depth_maps, for a normal pair and for an OCCLUDED RIGHT camera: detect_camera_occlusion must say 'right'
    and the blend takes the MiDaS-base branch (fused_depth_map.py:1665-1678).

A 1080p float map is 8 MB, so the float stages are stored as every 8th pixel in both directions plus a SHA-256 of the
exact integer-valued stages; the two final uint8 maps are stored whole (they compress well).  Build container only
(needs /root/reference and cv2):  python tests/golden/make_golden_fusion_1080p.py
"""
import contextlib
import hashlib
import io
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
sys.path.insert(0, ROOT)
from tests.golden_cases import FUSION_1080P, fusion_case_inputs  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)


def main():
    sys.path.insert(0, "/root/reference")
    with contextlib.redirect_stdout(io.StringIO()):
        import fused_depth_map as ref
    c = FUSION_1080P
    lb, rb, midas, flow = fusion_case_inputs(c)
    out = {}
    with contextlib.redirect_stdout(io.StringIO()):
        norm, disp, _, conf = ref.create_depth_map_stereo_scaled(lb, rb, c["min_disp"], c["num_disp"], c["window"])
        cal = ref.calibrate_midas_to_stereo(midas, disp, conf)
        flow_n = ref.normalize_to_stereo_range(flow, disp)
        fused, _ = ref.fuse_depth_maps(norm, conf, cal, None, flow_n, True)
        occ = ref.detect_camera_occlusion(lb, rb)
        # right camera covered: the detector's verdict switches stereo off, the blend starts from the mono map
        lb2, rb2, _, _ = fusion_case_inputs(dict(c, occluded="right"))
        occ2 = ref.detect_camera_occlusion(lb2, rb2)
        fused_m, _ = ref.fuse_depth_maps(None, None, cal, None, flow_n, True, use_stereo=False)
    assert occ[0] == "none" and occ2[0] == "right", (occ, occ2)
    out["disp_sha"], out["norm_sha"], out["conf_sha"] = sha(disp), sha(norm.astype(np.uint8)), sha(conf.astype(np.uint8))
    out["disp_valid"] = np.int64((disp > 0).sum())
    out["cal_s8"], out["flown_s8"] = cal[::8, ::8].copy(), flow_n[::8, ::8].copy()
    out["fused"], out["fused_m"] = fused, fused_m
    out["occ"] = np.array([occ[1], occ[2], occ2[1], occ2[2]], np.float64)
    path = os.path.join(HERE, "golden_fusion_1080p.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes; valid", int(out["disp_valid"]), "occ", occ, occ2)


if __name__ == "__main__":
    main()
