"""CPU: the oracle (oracle/*.c) against the cv2-made golden vectors, and against live cv2
where it is importable.  This is what pins the oracle (prompt section 3)."""
import numpy as np
import pytest

import oracle
from stereovision_b200.synth import make_pair, to_bgr
from tests.golden_cases import SGBM_CASES, SGBM_COLOR_CASES, color_pair


@pytest.mark.parametrize("i", range(len(SGBM_CASES)))
def test_sgbm_oracle_matches_golden(golden, i):
    c = SGBM_CASES[i]
    l, r, _ = make_pair(c["H"], c["W"], max(c["prm"]["numDisparities"], 16), seed=c["seed"])
    got = oracle.sgbm_compute(l, r, **c["prm"])
    ref = golden[f"sgbm_{i}"]
    assert got.dtype == np.int16 and got.shape == ref.shape
    assert int((got != ref).sum()) == 0          # bit-exact


@pytest.mark.parametrize("i", range(len(SGBM_COLOR_CASES)))
def test_sgbm_color_oracle_matches_golden(golden, i):
    c = SGBM_COLOR_CASES[i]
    l, r = color_pair(c)
    got = oracle.sgbm_compute(l, r, **c["prm"])
    assert got.shape == l.shape[:2] and int((got != golden[f"sgbm_color_{i}"]).sum()) == 0


def test_remap_gray_oracle_matches_golden(golden):
    l, _, _ = make_pair(120, 160, 16, seed=7)
    bgr = to_bgr(l)
    m1, m2 = golden["remap_map1"], golden["remap_map2"]
    assert np.array_equal(oracle.remap(bgr, m1, m2), golden["remap_bgr"])
    assert np.array_equal(oracle.remap(l, m1, m2), golden["remap_gray"])
    assert np.array_equal(oracle.bgr2gray(bgr), golden["gray"])


def test_median_speckle_oracle_matches_golden(golden):
    d16 = golden["med3_in"]
    assert np.array_equal(oracle.median3_s16(d16), golden["med3"])
    assert np.array_equal(oracle.median5_f32(d16.astype(np.float32) / 16.0), golden["med5"])
    prm = SGBM_CASES[0]["prm"]
    assert np.array_equal(oracle.filter_speckles(d16, (prm["minDisparity"] - 1) * 16, 40, 16), golden["speckle"])


def test_reproject_oracle_matches_golden(golden):
    d16 = golden["med3_in"]
    Q = golden["Q2"]
    for got, ref in ((oracle.reproject_image_to_3d(d16, Q), golden["reproj_s16"]),
                     (oracle.reproject_image_to_3d(d16.astype(np.float32) / 16.0, Q, True), golden["reproj_f32"])):
        fin = np.isfinite(ref)
        assert np.array_equal(fin, np.isfinite(got))
        np.testing.assert_allclose(got[fin], ref[fin], rtol=1e-4, atol=1e-6)   # north_star: 1e-4 relative


def test_sgbm_oracle_vs_live_cv2_random(have_cv2):
    if not have_cv2:
        pytest.skip("cv2 not importable here")
    import cv2
    rng = np.random.default_rng(5)
    for it in range(12):
        D = int(rng.choice([16, 32, 48]))
        bs = int(rng.choice([3, 5, 7]))
        mode = int(rng.choice([0, 1, 2, 3]))
        minD = 0 if mode == 2 else int(rng.choice([0, -8, 4]))
        H, W = int(rng.integers(40, 80)), int(rng.integers(D + 40, D + 120))
        prm = dict(minDisparity=minD, numDisparities=D, blockSize=bs, P1=8 * bs * bs, P2=32 * bs * bs,
                   disp12MaxDiff=int(rng.choice([1, 2])), preFilterCap=int(rng.choice([31, 63])),
                   uniquenessRatio=int(rng.choice([0, 10, 15])), speckleWindowSize=int(rng.choice([0, 60])),
                   speckleRange=2, mode=mode)
        l, r, _ = make_pair(H, W, D, seed=500 + it)
        assert np.array_equal(cv2.StereoSGBM_create(**prm).compute(l, r), oracle.sgbm_compute(l, r, **prm)), prm


def _bm_kwargs(c):
    kw = dict(numDisparities=c["numDisparities"], blockSize=c["blockSize"])
    m = {"MinDisparity": "minDisparity", "UniquenessRatio": "uniquenessRatio", "TextureThreshold": "textureThreshold",
         "SpeckleWindowSize": "speckleWindowSize", "SpeckleRange": "speckleRange", "PreFilterCap": "preFilterCap", "Disp12MaxDiff": "disp12MaxDiff",
         "PreFilterType": "preFilterType", "PreFilterSize": "preFilterSize"}
    for k, v in c.get("set", {}).items():
        kw[m[k]] = v
    return kw


def test_bm_oracle_matches_golden(golden):
    from tests.golden_cases import BM_CASES
    for i, c in enumerate(BM_CASES):
        l, r, _ = make_pair(c["H"], c["W"], max(c["numDisparities"], 16), seed=c["seed"], noise=c.get("noise", 6))
        got = oracle.bm_compute(l, r, **_bm_kwargs(c))
        assert np.array_equal(got, golden[f"bm_{i}"]), i


def test_validate_disparity_oracle_matches_golden(golden):
    d = golden["validate_in"]
    for k, (cost, d12) in enumerate(((golden["validate_cost16"], 1), (golden["validate_cost32"], 0),
                                     (golden["validate_cost16"], 4))):
        assert np.array_equal(oracle.validate_disparity(d, cost, 0, 16, d12), golden[f"validate_out_{k}"]), k


MAP_TAGS = {"l5": ("maps_R1", "maps_P1"), "r5": ("maps_R2", "maps_P2"), "l8": ("maps_R1", "maps_P1"),
            "l12": ("maps_R1", "maps_P1"), "l0": (None, "maps_K")}


def map_case(golden, tag):
    """(K, dist, R, P, size) of a golden initUndistortRectifyMap case; K of the right camera equals the left one
    in the synthetic calibration (tests/golden_cases.synth_calibration)."""
    rk, pk = MAP_TAGS[tag]
    dist = golden[f"maps_{tag}_dist"]
    return (golden["maps_K"], dist if dist.size else None, None if rk is None else golden[rk], golden[pk], (200, 140))


@pytest.mark.parametrize("tag", sorted(MAP_TAGS))
def test_undistort_rectify_map_oracle_matches_golden(golden, tag):
    from oracle.maps import init_undistort_rectify_map
    m1, m2 = init_undistort_rectify_map(*map_case(golden, tag))
    assert m1.dtype == np.int16 and m2.dtype == np.uint16
    assert np.array_equal(m1, golden[f"maps_{tag}_m1"]) and np.array_equal(m2, golden[f"maps_{tag}_m2"])


def test_resize_oracle_matches_golden(golden):
    """cv2.resize(INTER_LINEAR): u8 bit-exact, float32 within 1e-6 relative (fused_depth_map.py:473-476, 1214-1217)."""
    from tests.golden_cases import RESIZE_CASES, resize_case_input
    for i, c in enumerate(RESIZE_CASES):
        got = oracle.resize(resize_case_input(c), c["dsize"])
        ref = golden[f"resize_{i}"]
        assert got.shape == ref.shape and got.dtype == ref.dtype, i
        if c["ch"] == "f32":
            assert float(np.abs(got - ref).max()) <= 1e-6 * float(np.abs(ref).max()), i
        else:
            assert np.array_equal(got, ref), i
