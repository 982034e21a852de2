"""GPU: the per-frame functions of depth_map.py re-typed with `cv2.` -> `svb.` exactly as
INTEGRATION.md section 2 prescribes, against outputs of the UNMODIFIED reference functions
(tests/golden/golden_dropin.npz, made by tests/golden/make_golden_dropin.py from /root/reference).
"""
import os

import numpy as np
import pytest

import stereovision_b200 as svb
from stereovision_b200.synth import make_pair, to_bgr
from tests.conftest import ROOT
from tests.golden_cases import DROPIN_CASE

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gd():
    return np.load(os.path.join(ROOT, "tests", "golden", "golden_dropin.npz"))


def apply_stereo_rectification(left_img, right_img, stereo_calib):
    """depth_map.py:779-834 with the binding of INTEGRATION.md."""
    left_rectified = svb.remap(left_img, stereo_calib["left_map1"], stereo_calib["left_map2"], svb.INTER_LINEAR)
    right_rectified = svb.remap(right_img, stereo_calib["right_map1"], stereo_calib["right_map2"], svb.INTER_LINEAR)
    return left_rectified, right_rectified


def create_depth_map(left_img, right_img, MIN_DISP, NUM_DISP, WINDOW_SIZE, min_depth=0.3, max_depth=2.0):
    """depth_map.py:837-939 (without the colormap) with the binding of INTEGRATION.md."""
    gray_left = svb.cvtColor(left_img, svb.COLOR_BGR2GRAY) if left_img.ndim == 3 else left_img
    gray_right = svb.cvtColor(right_img, svb.COLOR_BGR2GRAY) if right_img.ndim == 3 else right_img
    stereo = svb.StereoSGBM_create(
        minDisparity=MIN_DISP, numDisparities=NUM_DISP, blockSize=WINDOW_SIZE,
        P1=8 * 3 * WINDOW_SIZE ** 2, P2=32 * 3 * WINDOW_SIZE ** 2, disp12MaxDiff=1, uniquenessRatio=10,
        speckleWindowSize=100, speckleRange=32, preFilterCap=63, mode=svb.STEREO_SGBM_MODE_SGBM_3WAY)
    disparity = stereo.compute(gray_left, gray_right).astype(np.float32) / 16.0
    disparity = svb.medianBlur(disparity, 5)
    fx = 700
    BASELINE = 0.08
    depth = (fx * BASELINE) / (disparity + 1e-6)
    depth_clipped = np.clip(depth, min_depth, max_depth)
    valid_mask = (disparity > MIN_DISP) & (depth_clipped >= min_depth) & (depth_clipped <= max_depth)
    return np.where(valid_mask, depth_clipped, 0), disparity


def _pair():
    c = DROPIN_CASE
    l, r, _ = make_pair(c["H"], c["W"], 64, seed=c["seed"])
    return to_bgr(l), to_bgr(r)


def test_rectification_bit_exact(gd):
    lb, rb = _pair()
    calib = {k: gd[k] for k in ("left_map1", "left_map2", "right_map1", "right_map2")}
    lr, rr = apply_stereo_rectification(lb, rb, calib)
    assert np.array_equal(lr, gd["left_rect"]) and np.array_equal(rr, gd["right_rect"])


def test_create_depth_map_matches_reference(gd):
    minD, D, ws = (int(v) for v in gd["params"])
    depth, disp = create_depth_map(gd["left_rect"], gd["right_rect"], minD, D, ws)
    assert np.array_equal(disp, gd["disparity"])                         # s16/16 + exact median: bit-exact
    np.testing.assert_allclose(depth, gd["depth"], rtol=1e-4, atol=0)    # north_star tolerance
    lb, rb = _pair()
    depth, disp = create_depth_map(lb, rb, minD, D, ws, min_depth=0.2, max_depth=40.0)
    assert np.array_equal(disp, gd["disparity_raw"])
    np.testing.assert_allclose(depth, gd["depth_raw"], rtol=1e-4, atol=0)
    assert float((depth > 0).mean()) > 0.2                               # the case exercises the valid branch


def test_fused_depth_kernel_matches_reference(gd):
    """The same tail as ONE kernel (svb.disparity_to_depth, include/svb.h svb_disp_to_depth_dev)."""
    minD, D, ws = (int(v) for v in gd["params"])
    gl = svb.cvtColor(gd["left_rect"], svb.COLOR_BGR2GRAY)
    gr = svb.cvtColor(gd["right_rect"], svb.COLOR_BGR2GRAY)
    d16 = svb.StereoSGBM_create(minDisparity=minD, numDisparities=D, blockSize=ws, P1=8 * 3 * ws ** 2,
                                P2=32 * 3 * ws ** 2, disp12MaxDiff=1, uniquenessRatio=10, speckleWindowSize=100,
                                speckleRange=32, preFilterCap=63, mode=svb.STEREO_SGBM_MODE_SGBM_3WAY).compute(gl, gr)
    depth, disp = svb.disparity_to_depth(d16, fx=700.0, baseline=0.08, min_disp=minD, min_depth=0.3, max_depth=2.0)
    assert np.array_equal(disp, gd["disparity"])
    np.testing.assert_allclose(depth, gd["depth"], rtol=1e-4, atol=0)
