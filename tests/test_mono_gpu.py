"""GPU: SURVEY 8(f) N3 (csrc/fuse.cu: bicubic upsample + normalise, contrast confidence, flow residual depth,
float32 bilateral filter) through stereovision_b200.mono, against oracle/mono.py and the golden vectors the
reference's own estimator classes produced."""
import os

import numpy as np
import pytest

import oracle
from oracle import fusion as F
from oracle import mono as M
import stereovision_b200 as svb
from tests.conftest import ROOT
from tests.golden_cases import MONO_CASES, mono_case_inputs
from tests.test_mono_oracle import close

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gm():
    return np.load(os.path.join(ROOT, "tests", "golden", "golden_mono.npz"))


@pytest.mark.parametrize("i", range(len(MONO_CASES)))
def test_midas_postprocess_matches_reference(gm, i):
    frame, pred = mono_case_inputs(MONO_CASES[i])
    depth, conf = svb.midas_postprocess(pred, frame)
    assert depth.dtype == np.float32 and conf.dtype == np.uint8 and depth.shape == frame.shape[:2]
    assert close(depth, gm[f"midas_depth_{i}"], 255.0)
    assert np.array_equal(conf, gm[f"midas_conf_{i}"])
    od, oc = M.midas_postprocess(pred, oracle.bgr2gray(frame))
    assert close(depth, od, 255.0) and np.array_equal(conf, oc)


@pytest.mark.parametrize("i", [i for i, c in enumerate(MONO_CASES) if not c.get("flat")])
def test_flow_residual_depth_matches_reference(gm, i):
    stable = None
    for k in range(3):
        sm, stable, cur = svb.flow_residual_depth(gm[f"flow_{i}_{k}"], gm[f"homography_{i}_{k}"], stable,
                                                  bool(gm[f"flow_moving_{i}_{k}"]))
        assert close(stable, gm[f"flow_stable_{i}_{k}"])
        assert close(sm, gm[f"flow_smoothed_{i}_{k}"])
        stable = gm[f"flow_stable_{i}_{k}"]                  # feed the reference's state: errors do not chain


def test_flow_residual_depth_static_alpha_and_batch(gm):
    rng = np.random.default_rng(9)
    flow = np.stack([gm["flow_0_0"], gm["flow_0_1"]])
    hm = np.stack([gm["homography_0_0"], gm["homography_0_1"]])
    prev = rng.uniform(0.1, 2.0, flow.shape[:3]).astype(np.float32)
    sm, st, cur = svb.flow_residual_depth(flow, hm, prev, camera_moving=False)
    for b in range(2):
        osm, ost, ocur = M.flow_residual_depth(flow[b], hm[b], prev[b], camera_moving=False)
        assert close(cur[b], ocur) and close(st[b], ost) and close(sm[b], osm)
    # degenerate homography row (w == 0 everywhere): cv2.perspectiveTransform returns (0, 0)
    z = np.zeros((3, 3))
    _, _, cur0 = svb.flow_residual_depth(flow[0], z)
    assert close(cur0, M.flow_residual_depth(flow[0], z)[2])


def test_contrast_confidence_shapes_and_edges():
    rng = np.random.default_rng(3)
    for shape in [(1, 1), (2, 3), (5, 300), (67, 129)]:
        g = rng.integers(0, 256, shape, dtype=np.uint8)
        assert np.array_equal(svb.contrast_confidence(g), M.contrast_confidence(g)), shape
    flat = np.full((40, 50), 9, np.uint8)
    assert not svb.contrast_confidence(flat).any()
    batch = rng.integers(0, 256, (3, 33, 70), dtype=np.uint8)
    batch[1] = 128                                              # per-image min / max
    out = svb.contrast_confidence(batch)
    for b in range(3):
        assert np.array_equal(out[b], M.contrast_confidence(batch[b]))
    bgr = rng.integers(0, 256, (31, 45, 3), dtype=np.uint8)
    assert np.array_equal(svb.contrast_confidence(bgr), M.contrast_confidence(oracle.bgr2gray(bgr)))


def test_upsample_normalize_sizes():
    rng = np.random.default_rng(4)
    for (hi, wi), (ho, wo) in [((1, 1), (5, 7)), ((3, 2), (64, 300)), ((40, 50), (40, 50)), ((90, 70), (17, 23))]:
        p = rng.normal(0, 3, (hi, wi)).astype(np.float32) if hi * wi > 1 else np.full((1, 1), 0.5, np.float32)
        got, want = svb.upsample_normalize(p, (ho, wo)), M.normalize_0_255(M.resize_bicubic(p, ho, wo))
        assert close(got, want, 255.0), ((hi, wi), (ho, wo), float(np.abs(got - want).max()))
    p = rng.normal(0, 3, (2, 12, 16)).astype(np.float32)
    out = svb.upsample_normalize(p, (30, 44))
    for b in range(2):
        assert close(out[b], M.normalize_0_255(M.resize_bicubic(p[b], 30, 44)), 255.0)


def test_bilateral_filter_f32(gm):
    gf = np.load(os.path.join(ROOT, "tests", "golden", "golden_fusion.npz"))
    from tests.golden_cases import FUSION_CASES, fusion_case_inputs
    flow = fusion_case_inputs(FUSION_CASES[0])[3]
    assert close(svb.bilateralFilter(flow, 9, 75, 75), gf["bilat9"])          # cv2's own output
    rng = np.random.default_rng(8)
    a = rng.uniform(0, 3, (37, 53)).astype(np.float32)                          # max < 10: still filtered
    for d, sc, ss in [(5, 0.5, 2.0), (9, 75, 75), (3, 1.0, 0.7)]:
        assert close(svb.bilateralFilter(a, d, sc, ss), F.bilateral_filter(a, d, sc, ss))
    flat = np.full((20, 30), 2.5, np.float32)
    assert np.array_equal(svb.bilateralFilter(flat, 9, 75, 75), flat)
    with pytest.raises(NotImplementedError):                     # d > 15 is outside the reference's use
        svb.bilateralFilter(a, 31, 75, 75)
