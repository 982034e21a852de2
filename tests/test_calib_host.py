"""CPU: the host side of the calibration-file path (stereovision_b200/calib.py: NumPy stereoRectify, Rodrigues,
the pickle reader) against goldens produced by the reference's own loaders (tests/golden/make_golden_calib.py:
depth_map.load_stereo_calibration, fused_depth_map.load_stereo_calibration_with_scaling, unmodified).  The maps
themselves are made on the device (tests/test_calib_gpu.py); here the oracle's map restatement stands in for the
kernel, which shows that the host algebra is precise enough for bit-identical maps."""
import os
import pickle

import numpy as np
import pytest

import oracle.maps as omaps
from stereovision_b200 import calib
from stereovision_b200.synth import synthetic_calibration, write_calibration_pkl
from tests.golden_cases import CALIB_CASE

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(ROOT, "tests", "golden", "golden_calib.npz"))


def _rectify(scale, flags=calib.CALIB_ZERO_DISPARITY):
    d = synthetic_calibration(CALIB_CASE["W"], CALIB_CASE["H"])
    Kl, Kr = d["mtx_left"].copy(), d["mtx_right"].copy()
    for K in (Kl, Kr):
        K[0, 0] *= scale; K[1, 1] *= scale; K[0, 2] *= scale; K[1, 2] *= scale
    size = (int(CALIB_CASE["W"] * scale), int(CALIB_CASE["H"] * scale))
    return d, Kl, Kr, size, calib.stereoRectify(Kl, d["dist_left"], Kr, d["dist_right"], size, d["R"], d["T"], alpha=0, flags=flags)


@pytest.mark.parametrize("scale", CALIB_CASE["scales"])
def test_stereo_rectify_matches_the_reference_loader(gold, scale):
    tag = f"scale{int(round(scale * 100))}"
    d, Kl, Kr, size, (R1, R2, P1, P2, Q, roi1, roi2) = _rectify(scale)
    assert tuple(gold[f"{tag}_size"]) == size
    assert np.allclose(Q, gold[f"{tag}_Q"], rtol=1e-12, atol=1e-12)
    assert np.array_equal(np.array([roi1, roi2]), gold[f"{tag}_roi"])
    assert gold[f"{tag}_focal_length"] == Kl[0, 0] and np.array_equal(gold[f"{tag}_mtx_left"], Kl)
    # maps from this R / P through the oracle's initUndistortRectifyMap: bit-identical to the reference's
    for side, K, dist, R, P in (("left", Kl, d["dist_left"], R1, P1), ("right", Kr, d["dist_right"], R2, P2)):
        m1, m2 = omaps.init_undistort_rectify_map(K, dist, R, P, size)
        assert np.array_equal(m1, gold[f"{tag}_{side}_map1"]) and np.array_equal(m2, gold[f"{tag}_{side}_map2"]), side


def test_plain_loader_golden_equals_scale_1(gold):
    """depth_map.load_stereo_calibration passes no flags: cv2's default is CALIB_ZERO_DISPARITY too."""
    assert np.array_equal(gold["plain_Q"], gold["scale100_Q"])
    assert np.array_equal(gold["plain_left_map1"], gold["scale100_left_map1"])
    _, _, _, _, out = _rectify(1.0)
    assert np.allclose(out[4], gold["plain_Q"], rtol=1e-12, atol=1e-12)


def test_rodrigues_round_trip_and_pickle_keys(tmp_path):
    rng = np.random.default_rng(3)
    for _ in range(20):
        v = rng.uniform(-1.5, 1.5, 3)
        R = calib.Rodrigues(v)
        assert np.allclose(R @ R.T, np.eye(3), atol=1e-14) and np.allclose(calib.Rodrigues(R).ravel(), v, atol=1e-12)
    assert np.array_equal(calib.Rodrigues(np.zeros(3)), np.eye(3))
    p = tmp_path / "output" / "stereo_calibration_data.pkl"
    write_calibration_pkl(str(p), 320, 240)
    with open(p, "rb") as f:
        d = pickle.load(f)
    assert set(d) == set(calib.CALIBRATION_KEYS) and tuple(d["img_size"]) == (320, 240)   # depth_map.py:613-619
    assert calib.load_stereo_calibration(str(tmp_path / "nope.pkl")) is None               # missing file -> None (depth_map.py:601-603)
