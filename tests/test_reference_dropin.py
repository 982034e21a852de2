"""GPU: the per-frame functions of depth_map.py re-typed with `cv2.` -> `hot.` exactly as INTEGRATION.md section 2
prescribes -- `hot` being the `_HotPath` alias of that section (svb where svb has the name, cv2 for everything else)
-- against outputs of the UNMODIFIED reference functions (tests/golden/golden_dropin.npz, made by
tests/golden/make_golden_dropin.py from /root/reference).  /root/reference does not exist on the GPU box and the
build container has no GPU, so the functions are re-typed here; the last test monkey-patches the imported reference
module itself and runs wherever both a GPU and /root/reference exist (skipped otherwise).
"""
import os

import numpy as np
import pytest

import stereovision_b200 as svb
from stereovision_b200.synth import make_pair, to_bgr
from tests.conftest import ROOT
from tests.golden_cases import DROPIN_CASE

pytestmark = pytest.mark.gpu

try:
    import cv2
except Exception:                                     # the alias then only ever resolves svb names
    cv2 = None


class _HotPath:                       # INTEGRATION.md section 2, verbatim: cv2 for capture/GUI, svb for the per-frame path
    def __getattr__(self, name):
        return getattr(svb, name) if hasattr(svb, name) else getattr(cv2, name)


hot = _HotPath()


@pytest.fixture(scope="module")
def gd():
    return np.load(os.path.join(ROOT, "tests", "golden", "golden_dropin.npz"))


def apply_stereo_rectification(left_img, right_img, stereo_calib):
    """depth_map.py:779-834 with the binding of INTEGRATION.md."""
    left_rectified = hot.remap(left_img, stereo_calib["left_map1"], stereo_calib["left_map2"], hot.INTER_LINEAR)
    right_rectified = hot.remap(right_img, stereo_calib["right_map1"], stereo_calib["right_map2"], hot.INTER_LINEAR)
    return left_rectified, right_rectified


def create_depth_map(left_img, right_img, MIN_DISP, NUM_DISP, WINDOW_SIZE, min_depth=0.3, max_depth=2.0):
    """depth_map.py:837-939 (without the colormap) with the binding of INTEGRATION.md."""
    gray_left = hot.cvtColor(left_img, hot.COLOR_BGR2GRAY) if left_img.ndim == 3 else left_img
    gray_right = hot.cvtColor(right_img, hot.COLOR_BGR2GRAY) if right_img.ndim == 3 else right_img
    stereo = hot.StereoSGBM_create(
        minDisparity=MIN_DISP, numDisparities=NUM_DISP, blockSize=WINDOW_SIZE,
        P1=8 * 3 * WINDOW_SIZE ** 2, P2=32 * 3 * WINDOW_SIZE ** 2, disp12MaxDiff=1, uniquenessRatio=10,
        speckleWindowSize=100, speckleRange=32, preFilterCap=63, mode=hot.STEREO_SGBM_MODE_SGBM_3WAY)
    disparity = stereo.compute(gray_left, gray_right).astype(np.float32) / 16.0
    disparity = hot.medianBlur(disparity, 5)
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


def test_hotpath_alias_resolves_svb_first_and_cv2_for_the_rest():
    assert hot.remap is svb.remap and hot.StereoSGBM_create is svb.StereoSGBM_create and hot.resize is svb.resize
    assert hot.STEREO_SGBM_MODE_SGBM_3WAY == 2 and hot.INTER_LINEAR == 1 and hot.COLOR_BGR2GRAY == 6
    if cv2 is not None:                               # GUI / capture names fall through to cv2
        assert hot.putText is cv2.putText and hot.FONT_HERSHEY_SIMPLEX == cv2.FONT_HERSHEY_SIMPLEX
        assert hot.INTER_LINEAR == cv2.INTER_LINEAR and hot.COLOR_BGR2GRAY == cv2.COLOR_BGR2GRAY


def test_monkeypatched_reference_module(gd, monkeypatch):
    """The reference's own depth_map.py with its module-level `cv2` replaced by the alias."""
    import contextlib
    import io
    import sys
    if not os.path.isdir("/root/reference") or cv2 is None:
        pytest.skip("needs /root/reference and cv2 next to the GPU")
    monkeypatch.syspath_prepend("/root/reference")
    with contextlib.redirect_stdout(io.StringIO()):
        import depth_map as ref
    monkeypatch.setattr(ref, "cv2", hot)
    lb, rb = _pair()
    calib = {k: gd[k] for k in ("left_map1", "left_map2", "right_map1", "right_map2")}
    with contextlib.redirect_stdout(io.StringIO()):
        lr, rr = ref.apply_stereo_rectification(lb, rb, calib)
        depth, disp, _ = ref.create_depth_map(lr, rr, None)
    assert np.array_equal(lr, gd["left_rect"]) and np.array_equal(rr, gd["right_rect"])
    assert np.array_equal(disp, gd["disparity"])
    np.testing.assert_allclose(depth, gd["depth"], rtol=1e-4, atol=0)


def test_frame_level_functions_match_the_reference(gd):
    """svb.frame.apply_stereo_rectification / create_depth_map: the reference's two per-frame functions with one upload
    and one download per frame (NumPy in / NumPy out, and tensors in / tensors out with device-resident maps)."""
    import torch
    minD, D, ws = (int(v) for v in gd["params"])
    lb, rb = _pair()
    H, W = lb.shape[:2]
    calib = {k: gd[k] for k in ("left_map1", "left_map2", "right_map1", "right_map2")}
    calib["img_size"] = (W, H)
    lr, rr = svb.frame.apply_stereo_rectification(lb, rb, calib)
    assert np.array_equal(lr, gd["left_rect"]) and np.array_equal(rr, gd["right_rect"])
    depth, disp, cmap = svb.frame.create_depth_map(lr, rr, None, min_disp=minD, num_disp=D, window_size=ws)
    assert depth.dtype == np.float32 and np.array_equal(disp, gd["disparity"])
    np.testing.assert_allclose(depth, gd["depth"], rtol=1e-4, atol=0)
    assert cmap.shape == (H, W, 3) and np.array_equal(cmap, svb.depth_colormap(disp, 700, 0.08, 0.3, 2.0))
    # a float64 fx from the calibration dict -> float64 depth like the reference (depth_map.py:917)
    c64 = {"calibration_data": {"camera_matrix": np.array([[700.0, 0, 0], [0, 700.0, 0], [0, 0, 1.0]])}}
    depth64, _, _ = svb.frame.create_depth_map(lr, rr, c64, min_disp=minD, num_disp=D, window_size=ws)
    assert depth64.dtype == np.float64 and np.allclose(depth64, gd["depth"], rtol=1e-4)
    # device-resident maps and frames: nothing touches the host
    dcal = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in calib.items() if k != "img_size"}
    dcal["img_size"] = (W, H)
    tl, tr = svb.frame.apply_stereo_rectification(torch.from_numpy(lb).cuda(), torch.from_numpy(rb).cuda(), dcal, as_tensor=True)
    td, tdisp, tc = svb.frame.create_depth_map(tl, tr, None, min_disp=minD, num_disp=D, window_size=ws, as_tensor=True)
    assert tdisp.is_cuda and np.array_equal(tdisp.cpu().numpy(), gd["disparity"]) and np.array_equal(tc.cpu().numpy(), cmap)
    # a frame of another size is resized to the calibration size first (depth_map.py:805-811)
    big = svb.resize(lb, (W + 40, H + 30))
    l2, r2 = svb.frame.apply_stereo_rectification(big, big, calib)
    assert l2.shape == lr.shape
    try:                                                      # ... bit for bit what the reference's two cv2 calls give
        import cv2
        small = cv2.resize(big, (W, H), interpolation=cv2.INTER_LINEAR)
        assert np.array_equal(l2, cv2.remap(small, gd["left_map1"], gd["left_map2"], cv2.INTER_LINEAR))
        assert np.array_equal(r2, cv2.remap(small, gd["right_map1"], gd["right_map2"], cv2.INTER_LINEAR))
    except ImportError:
        pass
    # the reference's error behaviour: zeros, never an exception (depth_map.py:941-946)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        z, zd, zc = svb.frame.create_depth_map(lr[:, :100], rr[:, :100], None, min_disp=minD, num_disp=D, window_size=ws)
    assert z.shape == (H, 100) and not z.any() and not zd.any() and zc.shape == (H, 100, 3)
