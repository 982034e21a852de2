"""CPU: oracle/fusion.py (NumPy restatement of the fusion chain) against the golden vectors that
tests/golden/make_golden_fusion.py produced by running the reference's own functions
(fused_depth_map.py imported from /root/reference).  This pins the fusion oracle."""
import os

import numpy as np
import pytest

import oracle
from oracle import fusion as F
from tests.conftest import ROOT
from tests.golden_cases import FUSION_CASES, fusion_case_inputs

RTOL = 1e-4          # north_star: float outputs within 1e-4 relative


@pytest.fixture(scope="module")
def gf():
    return np.load(os.path.join(ROOT, "tests", "golden", "golden_fusion.npz"))


def _close(a, b, scale=None):
    scale = float(np.abs(b).max()) if scale is None else scale
    return float(np.abs(a.astype(np.float64) - b.astype(np.float64)).max()) <= RTOL * max(scale, 1e-6)


def u8_close(got, ref, max_frac=0.15):
    """uint8 made by truncating a float chain: +-1 where the float lands on an integer boundary."""
    d = np.abs(got.astype(np.int32) - ref.astype(np.int32))
    return int(d.max()) <= 1 and float((d != 0).mean()) <= max_frac


@pytest.mark.parametrize("i", range(len(FUSION_CASES)))
def test_fusion_oracle_matches_reference(gf, i):
    c = FUSION_CASES[i]
    lb, rb, midas, flow = fusion_case_inputs(c)
    disp, conf, norm = gf[f"disp_{i}"], gf[f"conf_{i}"].astype(np.float32), gf[f"norm_{i}"].astype(np.float32)
    n2, c2 = F.normalize_disparity(disp, c["min_disp"], c["num_disp"])
    assert np.array_equal(n2, norm) and np.array_equal(c2, conf)            # integer-valued: exact
    cal = F.calibrate_midas_to_stereo(midas, disp, conf)
    assert _close(cal, gf[f"cal_{i}"])
    assert _close(F.calibrate_midas_to_stereo(midas, disp, np.zeros_like(conf)), gf[f"calfew_{i}"])
    assert _close(F.calibrate_midas_to_stereo(np.ascontiguousarray(midas[::2, ::3]), disp, conf), gf[f"calrs_{i}"])
    fn = F.normalize_to_stereo_range(flow, disp)
    assert _close(fn, gf[f"flown_{i}"])
    fu, _ = F.fuse_depth_maps(norm, conf, gf[f"cal_{i}"], None, gf[f"flown_{i}"], c["moving"], fusion_params=c["params"])
    assert u8_close(fu, gf[f"fused_{i}"])
    fu, _ = F.fuse_depth_maps(None, None, gf[f"cal_{i}"], None, gf[f"flown_{i}"], c["moving"], use_stereo=False,
                              fusion_params=c["params"])
    assert u8_close(fu, gf[f"fusedm_{i}"])
    fu, _ = F.fuse_depth_maps(None, None, None, None, gf[f"flown_{i}"], True, use_stereo=False, use_midas=False,
                              fusion_params=c["params"])
    assert u8_close(fu, gf[f"fusedf_{i}"])
    res, ls, rs = F.detect_camera_occlusion(oracle.bgr2gray(lb), oracle.bgr2gray(rb))
    assert {"none": 0, "left": 1, "right": 2, "both": 3}[res] == int(gf[f"occ_{i}"][0])
    assert abs(ls - gf[f"occ_{i}"][1]) < 1e-12 and abs(rs - gf[f"occ_{i}"][2]) < 1e-12


def test_filters_match_cv2_golden(gf):
    _, _, _, flow = fusion_case_inputs(FUSION_CASES[0])
    assert _close(F.gaussian_blur(flow), gf["gauss15"])
    assert _close(F.bilateral_filter(flow), gf["bilat9"])


def test_percentile_is_numpy_percentile():
    rng = np.random.default_rng(3)
    for it in range(60):
        a = rng.normal(0, 50, int(rng.integers(1, 4000))).astype(np.float32)
        if it % 3 == 0:
            a = np.round(a)
        for q in (0, 5, 10, 50, 90, 95, 100):
            assert F.percentile(a, q, True) == np.percentile(a, q)
        lo, hi = np.percentile(a, [10, 90])
        assert F.percentile(a, 10, False) == lo and F.percentile(a, 90, False) == hi


def test_fuse_returns_none_without_inputs():
    assert F.fuse_depth_maps(None, None, None, None, None, False) is None
    z = np.zeros((4, 4), np.float32)
    assert F.fuse_depth_maps(None, None, None, None, z, False, use_stereo=False, use_midas=False) is None   # flow needs motion
