"""GPU: BASELINE.json config 5 at its own size -- the fused_depth_map.py frame chain at 1920 x 1080 against goldens
made by the reference's own functions (tests/golden/make_golden_fusion_1080p.py): stereo stage (SGBM 3WAY D=256 ->
/16 -> median 5 -> normalise + confidence: integer-valued, compared by SHA-256), mono calibration and range
normalisation (float32, 1e-4 relative on every 8th pixel), the fused uint8 map (+-1 where the float chain lands on an
integer), and the occlusion fallback: right camera covered -> detector says 'right' -> MiDaS-base blend
(fused_depth_map.py:1665-1678)."""
import hashlib
import os

import numpy as np
import pytest

import stereovision_b200 as svb
from tests.conftest import ROOT
from tests.golden_cases import FUSION_1080P, fusion_case_inputs
from tests.test_fusion_oracle import _close, u8_close

pytestmark = pytest.mark.gpu


def _sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)


def test_config5_chain_at_1080p_matches_the_reference():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    g = np.load(os.path.join(ROOT, "tests", "golden", "golden_fusion_1080p.npz"))
    c = FUSION_1080P
    lb, rb, midas, flow = fusion_case_inputs(c)
    # create_depth_map_stereo_scaled (fused_depth_map.py:934-1029)
    gl, gr = svb.cvtColor(lb, svb.COLOR_BGR2GRAY), svb.cvtColor(rb, svb.COLOR_BGR2GRAY)
    w = c["window"]
    m = svb.StereoSGBM_create(minDisparity=c["min_disp"], numDisparities=c["num_disp"], blockSize=w, P1=8 * 3 * w ** 2,
                              P2=32 * 3 * w ** 2, disp12MaxDiff=1, uniquenessRatio=10, speckleWindowSize=100,
                              speckleRange=32, preFilterCap=63, mode=svb.STEREO_SGBM_MODE_SGBM_3WAY)
    disp = svb.medianBlur(m.compute(gl, gr).astype(np.float32) / 16.0, 5)
    assert int((disp > 0).sum()) == int(g["disp_valid"]) and np.array_equal(_sha(disp), g["disp_sha"])
    norm, conf = svb.normalize_disparity(disp, c["min_disp"], c["num_disp"])
    assert np.array_equal(_sha(norm.astype(np.uint8)), g["norm_sha"]) and np.array_equal(_sha(conf.astype(np.uint8)), g["conf_sha"])
    # calibrate_midas_to_stereo / normalize_to_stereo_range (:1169-1257, :1503-1554)
    cal = svb.calibrate_midas_to_stereo(midas, disp, conf)
    flow_n = svb.normalize_to_stereo_range(flow, disp)
    assert _close(cal[::8, ::8], g["cal_s8"]) and _close(flow_n[::8, ::8], g["flown_s8"])
    # fuse_depth_maps (:1560-1698), stereo branch
    fused = svb.fuse_depth_maps(norm, conf, cal, None, flow_n, True)
    assert fused.shape == (c["H"], c["W"]) and u8_close(fused, g["fused"])
    # occlusion fallback: the detector (:131-301) on the covered right camera, then the MiDaS-base branch
    res, ls, rs = svb.detect_camera_occlusion(lb, rb)
    assert res == "none" and np.allclose([ls, rs], g["occ"][:2], atol=1e-9)
    lb2, rb2, _, _ = fusion_case_inputs(dict(c, occluded="right"))
    res2, ls2, rs2 = svb.detect_camera_occlusion(lb2, rb2)
    assert res2 == "right" and np.allclose([ls2, rs2], g["occ"][2:], atol=1e-9)
    mon = svb.OcclusionMonitor()
    for k in range(10):
        use_stereo = mon.update(k, lb2, rb2)
    assert not use_stereo and mon.state == "right"
    fused_m = svb.fuse_depth_maps(None, None, cal, None, flow_n, True, use_stereo=use_stereo)
    assert u8_close(fused_m, g["fused_m"])
    # device-resident: the same chain on tensors gives the same bytes
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    fd = svb.fuse_depth_maps(t(norm), t(conf), t(cal), None, t(flow_n), True)
    assert np.array_equal(fd.cpu().numpy(), fused)
