"""CPU: oracle/mono.py against tests/golden/golden_mono.npz, which the reference's own
DepthEstimatorMidas / OpticalFlowDepthEstimator produced (tests/golden/make_golden_mono.py)."""
import os

import numpy as np
import pytest

import oracle
from oracle import mono as M
from tests.conftest import ROOT
from tests.golden_cases import MONO_CASES, mono_case_inputs

RTOL = 1e-4            # BASELINE.json north_star: <= 1e-4 relative for floating-point outputs


def close(a, b, scale=None):
    scale = max(float(np.abs(b).max()), 1e-6) if scale is None else scale
    return float(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64)).max()) <= RTOL * scale


@pytest.fixture(scope="module")
def gm():
    return np.load(os.path.join(ROOT, "tests", "golden", "golden_mono.npz"))


@pytest.mark.parametrize("i", range(len(MONO_CASES)))
def test_midas_postprocess_matches_reference(gm, i):
    frame, pred = mono_case_inputs(MONO_CASES[i])
    depth, conf = M.midas_postprocess(pred, oracle.bgr2gray(frame))
    assert depth.dtype == np.float32 and conf.dtype == np.uint8
    assert close(depth, gm[f"midas_depth_{i}"], 255.0)
    assert np.array_equal(conf, gm[f"midas_conf_{i}"])              # integer output: exact


@pytest.mark.parametrize("i", [i for i, c in enumerate(MONO_CASES) if not c.get("flat")])
def test_flow_residual_depth_matches_reference(gm, i):
    stable = None
    for k in range(3):
        sm, stable, cur = M.flow_residual_depth(gm[f"flow_{i}_{k}"], gm[f"homography_{i}_{k}"], stable,
                                                bool(gm[f"flow_moving_{i}_{k}"]))
        assert np.array_equal(stable, gm[f"flow_stable_{i}_{k}"])   # elementwise float32: bit-equal on this libm
        assert close(sm, gm[f"flow_smoothed_{i}_{k}"])


def test_u8_gaussian_and_minmax_edge_cases():
    rng = np.random.default_rng(5)
    one = np.full((7, 9), 200, np.uint8)
    assert np.array_equal(M.gaussian_blur5_u8(one), one)
    assert not M.contrast_confidence(one).any()                      # flat image: scale 0 -> zeros
    tiny = rng.integers(0, 256, (2, 3), dtype=np.uint8)              # smaller than the kernel: reflect101 twice
    assert M.gaussian_blur5_u8(tiny).shape == tiny.shape
    c = np.array([[3, 4, 9]], np.uint8)
    assert np.array_equal(M.normalize_minmax_u8(c), np.array([[0, 0, 1]], np.uint8))   # 1/6 -> 0, ties to even
    assert np.array_equal(M.normalize_minmax_u8(np.array([[0, 5, 10]], np.uint8)), np.array([[0, 0, 1]], np.uint8))
