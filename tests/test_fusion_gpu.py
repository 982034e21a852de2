"""GPU: the fusion chain (csrc/fuse.cu) through the reference-named Python surface, against the
oracle (oracle/fusion.py) and the golden vectors made by the reference's own functions."""
import os

import numpy as np
import pytest

import oracle
from oracle import fusion as F
import stereovision_b200 as svb
from stereovision_b200.synth import make_pair, make_aux_depths
from tests.conftest import ROOT
from tests.golden_cases import FUSION_CASES, fusion_case_inputs
from tests.test_fusion_oracle import RTOL, _close, u8_close

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gf():
    return np.load(os.path.join(ROOT, "tests", "golden", "golden_fusion.npz"))


@pytest.mark.parametrize("i", range(len(FUSION_CASES)))
def test_stereo_stage_matches_reference(gf, i):
    """create_depth_map_stereo_scaled (fused_depth_map.py:934-1029) end to end: gray -> SGBM 3WAY ->
    /16 -> medianBlur 5 -> normalise + confidence; all integer-valued, so exact."""
    c = FUSION_CASES[i]
    lb, rb, _, _ = fusion_case_inputs(c)
    gl, gr = svb.cvtColor(lb, svb.COLOR_BGR2GRAY), svb.cvtColor(rb, svb.COLOR_BGR2GRAY)
    w = c["window"]
    m = svb.StereoSGBM_create(minDisparity=c["min_disp"], numDisparities=c["num_disp"], blockSize=w,
                              P1=8 * 3 * w ** 2, P2=32 * 3 * w ** 2, disp12MaxDiff=1, uniquenessRatio=10,
                              speckleWindowSize=100, speckleRange=32, preFilterCap=63,
                              mode=svb.STEREO_SGBM_MODE_SGBM_3WAY)
    disp = svb.medianBlur(m.compute(gl, gr).astype(np.float32) / 16.0, 5)
    assert np.array_equal(disp, gf[f"disp_{i}"])
    norm, conf = svb.normalize_disparity(disp, c["min_disp"], c["num_disp"])
    assert np.array_equal(norm, gf[f"norm_{i}"].astype(np.float32))
    assert np.array_equal(conf, gf[f"conf_{i}"].astype(np.float32))


@pytest.mark.parametrize("i", range(len(FUSION_CASES)))
def test_calibration_and_range_match_reference(gf, i):
    c = FUSION_CASES[i]
    _, _, midas, flow = fusion_case_inputs(c)
    disp, conf = gf[f"disp_{i}"], gf[f"conf_{i}"].astype(np.float32)
    assert _close(svb.calibrate_midas_to_stereo(midas, disp, conf), gf[f"cal_{i}"])
    assert _close(svb.calibrate_midas_to_stereo(midas, disp, np.zeros_like(conf)), gf[f"calfew_{i}"])
    # a mono map of another shape is resized to the stereo shape on the device first (fused_depth_map.py:1214-1217)
    small = np.ascontiguousarray(midas[::2, ::3])
    assert _close(svb.calibrate_midas_to_stereo(small, disp, conf), gf[f"calrs_{i}"])
    import torch
    dev = svb.calibrate_midas_to_stereo(torch.from_numpy(small).cuda(), torch.from_numpy(disp).cuda(), torch.from_numpy(conf).cuda())
    assert _close(dev.cpu().numpy(), gf[f"calrs_{i}"])
    assert _close(svb.normalize_to_stereo_range(flow, disp), gf[f"flown_{i}"])


@pytest.mark.parametrize("i", range(len(FUSION_CASES)))
def test_fuse_matches_reference_and_oracle(gf, i):
    c = FUSION_CASES[i]
    norm, conf = gf[f"norm_{i}"].astype(np.float32), gf[f"conf_{i}"].astype(np.float32)
    cal, fn = gf[f"cal_{i}"], gf[f"flown_{i}"]
    for args, kw, key in (((norm, conf, cal, None, fn, c["moving"]), {}, "fused"),
                          ((None, None, cal, None, fn, c["moving"]), dict(use_stereo=False), "fusedm"),
                          ((None, None, None, None, fn, True), dict(use_stereo=False, use_midas=False), "fusedf")):
        u8, f = svb.fuse_depth_maps(*args, fusion_params=c["params"], return_float=True, **kw)
        assert u8.dtype == np.uint8 and u8_close(u8, gf[f"{key}_{i}"])
        o8, of = F.fuse_depth_maps(*args, fusion_params=c["params"], **kw)
        assert _close(f, of, scale=255.0)
        assert u8_close(u8, o8)


def test_percentiles_exact_order_statistics():
    """np.percentile semantics: scalar q in float32, list q in float64, masks, ties, tiny inputs."""
    rng = np.random.default_rng(9)
    for n, H, W in ((1, 1, 1), (2, 1, 2), (7, 1, 7), (640, 20, 32), (99 * 50, 99, 50), (200 * 301, 200, 301)):
        a = rng.normal(0, 40, (3, H, W)).astype(np.float32)
        a[1] = np.round(a[1] / 8) * 8                       # heavy ties
        a[2] = np.abs(a[2]) * (rng.random((H, W)) > 0.5)    # many zeros
        got, cnt = svb.percentiles(a)
        for b in range(3):
            assert cnt[b] == n
            assert got[b, 0] == np.percentile(a[b], 5) and got[b, 1] == np.percentile(a[b], 95)
        got, cnt = svb.percentiles(a, mask_mode=2)           # values > 0
        for b in range(3):
            sel = a[b][a[b] > 0]
            assert cnt[b] == sel.size
            if sel.size:
                assert got[b, 0] == np.percentile(sel, 5) and got[b, 1] == np.percentile(sel, 95)
            else:
                assert np.isnan(got[b]).all()
        mask = rng.random((3, H, W)).astype(np.float32)
        got, cnt = svb.percentiles(a, mask=mask, mask_thr=0.7, mask_mode=3)
        for b in range(3):
            sel = a[b][mask[b] > 0.7]
            if sel.size < 100:
                assert cnt[b] == n
                assert got[b, 0] == np.percentile(a[b], 5) and got[b, 1] == np.percentile(a[b], 95)
            else:
                lo, hi = np.percentile(sel, [10, 90])
                assert cnt[b] == sel.size and got[b, 0] == lo and got[b, 1] == hi


def test_range_degenerate_cases():
    z = np.zeros((40, 50), np.float32)
    d = np.full((40, 50), 7.0, np.float32)
    rng = np.random.default_rng(2)
    m = rng.normal(500, 80, (40, 50)).astype(np.float32)
    # constant map -> (smin + smax) / 2 ; no valid stereo -> 0..255 defaults
    assert np.array_equal(svb.normalize_to_stereo_range(d, z), F.normalize_to_stereo_range(d, z))
    assert _close(svb.normalize_to_stereo_range(m, z), F.normalize_to_stereo_range(m, z))
    s = np.abs(rng.normal(20, 10, (40, 50))).astype(np.float32)
    assert np.array_equal(svb.calibrate_midas_to_stereo(d, s, z), F.calibrate_midas_to_stereo(d, s, z))
    one = np.ones_like(z)
    assert _close(svb.calibrate_midas_to_stereo(d, s, one), F.calibrate_midas_to_stereo(d, s, one))   # scale = 1 branch


def test_fuse_batched_and_tensor_io():
    import torch
    H, W = 150, 210
    S, C, M, Fl = [], [], [], []
    for k in range(3):
        _, _, gt = make_pair(H, W, 64, seed=60 + k)
        midas, flow = make_aux_depths(gt, 60 + k)
        n, c = F.normalize_disparity(gt.astype(np.float32) - (k == 1) * 3, 0, 64)
        S.append(n); C.append(c); Fl.append(flow)
        M.append(F.calibrate_midas_to_stereo(midas, gt.astype(np.float32), c))
    S, C, M, Fl = (np.stack(x) for x in (S, C, M, Fl))
    got = svb.fuse_depth_maps(*(torch.from_numpy(x).cuda() for x in (S, C, M)), None, torch.from_numpy(Fl).cuda(), True)
    assert got.is_cuda and got.dtype == torch.uint8
    got = got.cpu().numpy()
    for k in range(3):
        ref, _ = F.fuse_depth_maps(S[k], C[k], M[k], None, Fl[k], True)
        assert u8_close(got[k], ref)
    # low values only: the bilateral stage is skipped (max <= 10)
    lo = np.full((H, W), 5.0, np.float32)
    u8, f = svb.fuse_depth_maps(lo, np.ones_like(lo), None, None, None, False, return_float=True)
    assert np.array_equal(f, np.full((H, W), np.float32(5.0) * np.float32(0.8)))
    assert svb.fuse_depth_maps(None, None, None, None, None, True) is None


@pytest.mark.parametrize("i", range(len(FUSION_CASES)))
def test_occlusion_detector_matches_reference(gf, i):
    c = FUSION_CASES[i]
    lb, rb, _, _ = fusion_case_inputs(c)
    res, ls, rs = svb.detect_camera_occlusion(lb, rb)
    assert {"none": 0, "left": 1, "right": 2, "both": 3}[res] == int(gf[f"occ_{i}"][0])
    assert abs(ls - gf[f"occ_{i}"][1]) < 1e-12 and abs(rs - gf[f"occ_{i}"][2]) < 1e-12
    gl = oracle.bgr2gray(lb)
    st = svb.occlusion_stats(gl)
    np.testing.assert_allclose(st, np.array(F.occlusion_stats(gl)), rtol=RTOL)


def test_occlusion_stats_shapes():
    rng = np.random.default_rng(4)
    for H, W in ((30, 40), (48, 48), (100, 130), (97, 200)):    # smaller than a block, exact, ragged
        g = rng.integers(0, 256, (2, H, W), dtype=np.uint8)
        g[1] = (g[1] // 32) + 100
        st = svb.occlusion_stats(g)
        for b in range(2):
            np.testing.assert_allclose(st[b], np.array(F.occlusion_stats(g[b])), rtol=RTOL, atol=1e-9)


def test_occlusion_monitor_replays_a_left_occluded_sequence():
    """fused_depth_map.py:2515-2571 with the real detector: the left camera is covered from frame 4 on -> five
    differing checks (every 2nd frame) later stereo is switched off; uncovered again -> switched back on."""
    from stereovision_b200.synth import make_pair, to_bgr
    l, r, _ = make_pair(240, 320, 32, seed=5)
    lb, rb = to_bgr(l), to_bgr(r)
    covered = np.full_like(lb, 22) + (lb // 64)
    mon = svb.OcclusionMonitor()
    hist = []
    for k in range(40):
        left = covered if 4 <= k < 24 else lb
        hist.append(mon.update(k, left, rb))
    assert mon.last_detected == "none"
    assert hist[:12] == [True] * 12 and hist[12:32] == [False] * 20 and hist[32:] == [True] * 8
