"""GPU: libsvb.so SGBM against the oracle on seeded inputs, against the cv2-made golden
vectors, stage by stage (C, S, raw), batched, and via size-independent properties at
BASELINE sizes.  Bit-exact: int16 disparities must be identical."""
import numpy as np
import pytest

import oracle
import stereovision_b200 as svb
from stereovision_b200.synth import make_pair, make_batch
from tests.golden_cases import SGBM_CASES, SGBM_COLOR_CASES, color_pair

pytestmark = pytest.mark.gpu


def _torch():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch


@pytest.mark.parametrize("i", range(len(SGBM_CASES)))
def test_sgbm_matches_golden_and_oracle(golden, i):
    _torch()
    c = SGBM_CASES[i]
    l, r, _ = make_pair(c["H"], c["W"], max(c["prm"]["numDisparities"], 16), seed=c["seed"])
    m = svb.StereoSGBM_create(**c["prm"])
    m.setDebugStages(True)
    got = m.compute(l, r)
    assert got.dtype == np.int16 and got.shape == l.shape
    ref, st = oracle.sgbm_compute(l, r, stages=True, **c["prm"])
    if not np.array_equal(got, ref):
        C, S, raw = m.debug_stages(*l.shape)
        keep = np.ones(l.shape[0], bool)
        msg = (f"C mismatches {int((C != st['C'].view(np.uint16)).sum())}, S mismatches {int((S != st['S']).sum())}, "
               f"raw mismatches {int((raw != st['raw']).sum())}, final {int((got != ref).sum())}")
        raise AssertionError(msg)
    assert np.array_equal(got, golden[f"sgbm_{i}"])


@pytest.mark.parametrize("i", range(len(SGBM_COLOR_CASES)))
def test_sgbm_color_matches_golden_and_oracle(golden, i):
    """3-channel input: [H,W,3] like cv2, [B,H,W,3] batched, host and device paths."""
    torch = _torch()
    c = SGBM_COLOR_CASES[i]
    l, r = color_pair(c)
    m = svb.StereoSGBM_create(**c["prm"])
    got = m.compute(l, r)
    assert got.dtype == np.int16 and got.shape == l.shape[:2]
    assert np.array_equal(got, oracle.sgbm_compute(l, r, **c["prm"]))
    assert np.array_equal(got, golden[f"sgbm_color_{i}"])
    L2, R2 = np.stack([l, l[::-1].copy()]), np.stack([r, r[::-1].copy()])
    dev = m.compute(torch.from_numpy(L2).cuda(), torch.from_numpy(R2).cuda()).cpu().numpy()
    assert dev.shape == L2.shape[:3] and np.array_equal(dev[0], got)
    assert np.array_equal(dev[1], oracle.sgbm_compute(L2[1], R2[1], **c["prm"]))
    with pytest.raises(ValueError):
        svb.StereoBM_create(16, 9).compute(l, r)


@pytest.mark.parametrize("mode", [0, 1, 2, 3])
def test_sgbm_stages_match_oracle(mode):
    _torch()
    H, W, D, bs = 50, 150, 64, 5
    l, r, _ = make_pair(H, W, D, seed=77)
    prm = dict(minDisparity=0, numDisparities=D, blockSize=bs, P1=8 * bs * bs, P2=32 * bs * bs, disp12MaxDiff=1,
               preFilterCap=63, uniquenessRatio=10, speckleWindowSize=0, speckleRange=0, mode=mode)
    m = svb.StereoSGBM_create(**prm)
    m.setDebugStages(True)
    got = m.compute(l, r)
    ref, st = oracle.sgbm_compute(l, r, stages=True, **prm)
    C, S, raw = m.debug_stages(H, W)
    assert np.array_equal(C, st["C"].view(np.uint16)), "cost volume"
    assert np.array_equal(S, st["S"]), "aggregated volume"
    assert np.array_equal(raw, st["raw"]), "raw disparity"
    assert np.array_equal(got, ref)


def test_sgbm_random_params_vs_oracle():
    _torch()
    rng = np.random.default_rng(123)
    for it in range(48):
        D = int(rng.choice([16, 32, 48, 64, 96, 128, 192]))
        bs = int(rng.choice([1, 3, 5, 7, 9, 11]))
        mode = int(rng.choice([0, 1, 2, 3]))
        minD = int(rng.choice([0, 0, -8, 4, 16]))
        if mode == 2 and bs == 1:
            bs = 3
        H = int(rng.integers(20, 70))
        W = int(rng.integers(D + abs(minD) + bs + 8, D + abs(minD) + 120))
        prm = dict(minDisparity=minD, numDisparities=D, blockSize=bs, P1=int(rng.choice([0, 8 * bs * bs])),
                   P2=int(rng.choice([0, 32 * bs * bs])), disp12MaxDiff=int(rng.choice([-1, 1, 3])),
                   preFilterCap=int(rng.choice([0, 31, 63])), uniquenessRatio=int(rng.choice([0, 5, 10, 30])),
                   speckleWindowSize=int(rng.choice([0, 40, 100])), speckleRange=int(rng.choice([1, 2, 32])), mode=mode)
        l, r, _ = make_pair(H, W, max(D, 16), seed=900 + it, noise=int(rng.choice([0, 6, 20])))
        if rng.random() < 0.25:
            l = rng.integers(0, 256, (H, W), dtype=np.uint8)
            r = rng.integers(0, 256, (H, W), dtype=np.uint8)
        got = svb.StereoSGBM_create(**prm).compute(l, r)
        ref = oracle.sgbm_compute(l, r, **prm)
        assert np.array_equal(got, ref), (H, W, prm, int((got != ref).sum()))


def test_sgbm_batch_equals_single_and_device_path():
    torch = _torch()
    B, H, W, D = 5, 48, 160, 64
    L, R = make_batch(B, H, W, D, seed0=40)
    prm = dict(minDisparity=0, numDisparities=D, blockSize=5, P1=200, P2=800, disp12MaxDiff=1, preFilterCap=63,
               uniquenessRatio=10, speckleWindowSize=100, speckleRange=32, mode=svb.STEREO_SGBM_MODE_SGBM)
    m = svb.StereoSGBM_create(**prm)
    host = m.compute(L, R)                                   # host path: H2D + compute + D2H inside the C call
    dev = m.compute(torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda())
    assert dev.is_cuda and dev.dtype == torch.int16
    assert np.array_equal(host, dev.cpu().numpy())
    for b in range(B):
        assert np.array_equal(host[b], m.compute(L[b], R[b]))
        if b < 2:
            assert np.array_equal(host[b], oracle.sgbm_compute(L[b], R[b], **prm))


def test_sgbm_chunked_workspace_gives_same_result():
    _torch()
    from stereovision_b200 import _lib
    import ctypes
    B, H, W, D = 6, 40, 140, 32
    L, R = make_batch(B, H, W, D, seed0=60)
    prm = (0, D, 5, 200, 800, 1, 63, 10, 50, 2, 1)
    lib = _lib.lib()
    outs = []
    for limit in (0, 1):       # 1 byte limit -> one image per chunk
        h = ctypes.c_void_p()
        _lib.check(lib.svb_sgbm_create(ctypes.byref(_lib.SgbmParams(*prm)), ctypes.byref(h)))
        lib.svb_sgbm_set_workspace_limit(h, limit)
        out = np.empty((B, H, W), np.int16)
        _lib.check(lib.svb_sgbm_compute_host(h, L.ctypes.data, R.ctypes.data, B, H, W, out.ctypes.data))
        lib.svb_sgbm_destroy(h)
        outs.append(out)
    assert np.array_equal(outs[0], outs[1])


def test_sgbm_errors_like_cv2():
    _torch()
    with pytest.raises(ValueError):      # width too small for the disparity range (cv2 asserts)
        svb.StereoSGBM_create(0, 64, 5).compute(np.zeros((20, 60), np.uint8), np.zeros((20, 60), np.uint8))
    with pytest.raises(NotImplementedError):      # cv2 accepts it; this engine says so instead of differing
        svb.StereoSGBM_create(0, 24, 3).compute(np.zeros((20, 60), np.uint8), np.zeros((20, 60), np.uint8))


@pytest.mark.parametrize("cfg", [
    dict(H=480, W=640, D=64, mode=0),          # BASELINE config 1 shape
    dict(H=720, W=1280, D=128, mode=0),        # config 3 shape
    dict(H=1080, W=1920, D=256, mode=1),       # config 4 shape
    dict(H=480, W=640, D=320, mode=2),         # the reference's own numDisparities, 3WAY
    dict(H=720, W=1280, D=128, mode=3),        # HH4
])
def test_sgbm_full_size_properties(cfg):
    """At BASELINE sizes the oracle is too slow; check size-independent properties:
    determinism, batch-order independence, valid range, and recovery of the known
    synthetic disparity on unoccluded pixels."""
    _torch()
    H, W, D = cfg["H"], cfg["W"], cfg["D"]
    prm = dict(minDisparity=0, numDisparities=D, blockSize=5, P1=200, P2=800, disp12MaxDiff=1, preFilterCap=63,
               uniquenessRatio=10, speckleWindowSize=100, speckleRange=32, mode=cfg["mode"])
    l0, r0, gt0 = make_pair(H, W, D, seed=1)
    l1, r1, _ = make_pair(H, W, D, seed=2)
    m = svb.StereoSGBM_create(**prm)
    a = m.compute(np.stack([l0, l1]), np.stack([r0, r1]))
    b = m.compute(np.stack([l1, l0]), np.stack([r1, r0]))
    assert np.array_equal(a[0], b[1]) and np.array_equal(a[1], b[0])
    assert a.min() >= -16 and a.max() <= (D - 1) * 16
    assert np.all(a[0][:, :D] == -16)                     # columns x < D are never matched (A.1)
    valid = a[0] >= 0
    assert valid[:, D:].mean() > 0.5                      # of the columns that can be matched at all
    err = np.abs(a[0][valid] / 16.0 - gt0[valid])
    assert np.mean(err <= 1.0) > 0.95


@pytest.mark.parametrize("cfg", [
    dict(H=64, W=200, D=64, mode=1, B=600),     # more bands than CTAs: every CTA takes several bands in turn
    dict(H=100, W=500, D=256, mode=1, B=100),   # the headline vector width (4 words per lane), carry across bands
    dict(H=64, W=200, D=64, mode=0, B=600),
    dict(H=70, W=420, D=320, mode=2, B=80),     # 3WAY stripes at the reference's own numDisparities
    dict(H=90, W=300, D=128, mode=3, B=200),    # HH4: both sweeps carry a vertical path across bands
])
def test_sgbm_many_pairs_persistent_ctas_bit_exact(cfg):
    """The sweep kernels are persistent: with more (pair, band) tickets than resident CTAs each CTA
    re-initialises its rings and counters between bands.  Every copy of the pair must equal the oracle."""
    torch = _torch()
    H, W, D, B = cfg["H"], cfg["W"], cfg["D"], cfg["B"]
    l, r, _ = make_pair(H, W, D, seed=3)
    bs = 5
    prm = dict(minDisparity=0, numDisparities=D, blockSize=bs, P1=8 * bs * bs, P2=32 * bs * bs, disp12MaxDiff=1,
               preFilterCap=63, uniquenessRatio=10, speckleWindowSize=100, speckleRange=32, mode=cfg["mode"])
    ref = torch.from_numpy(oracle.sgbm_compute(l, r, **prm)).cuda()
    L = torch.from_numpy(l).cuda().unsqueeze(0).repeat(B, 1, 1).contiguous()
    R = torch.from_numpy(r).cuda().unsqueeze(0).repeat(B, 1, 1).contiguous()
    out = svb.StereoSGBM_create(**prm).compute(L, R)
    assert int((out != ref.unsqueeze(0)).sum()) == 0


@pytest.mark.parametrize("mode", [0, 2])
def test_sgbm_single_live_frame_bit_exact(mode):
    """One 640x480 frame (the reference's camera loop): a single image takes the small-batch path (short bands in
    the sweep without inter-row paths, host-buffer pipeline with one sub-range)."""
    _torch()
    H, W, D, bs = 480, 640, 64, 5
    l, r, _ = make_pair(H, W, D, seed=12)
    prm = dict(minDisparity=0, numDisparities=D, blockSize=bs, P1=8 * bs * bs, P2=32 * bs * bs, disp12MaxDiff=1,
               preFilterCap=63, uniquenessRatio=10, speckleWindowSize=100, speckleRange=32, mode=mode)
    assert np.array_equal(svb.StereoSGBM_create(**prm).compute(l, r), oracle.sgbm_compute(l, r, **prm))


def test_sgbm_color_random_params_vs_oracle():
    """3-channel input over random parameters, modes and shapes (the oracle's colour path is pinned by live cv2
    and by the sgbm_color_* goldens)."""
    _torch()
    rng = np.random.default_rng(321)
    for it in range(12):
        D = int(rng.choice([16, 32, 64, 96, 160]))
        bs = int(rng.choice([1, 3, 5, 7]))
        mode = int(rng.choice([0, 1, 2, 3]))
        minD = 0 if mode == 2 else int(rng.choice([0, -8, 4]))
        if mode == 2 and bs == 1:
            bs = 3
        H = int(rng.integers(20, 60))
        W = int(rng.integers(D + abs(minD) + bs + 8, D + abs(minD) + 100))
        prm = dict(minDisparity=minD, numDisparities=D, blockSize=bs, P1=8 * 3 * bs * bs, P2=32 * 3 * bs * bs,
                   disp12MaxDiff=int(rng.choice([1, 2])), preFilterCap=int(rng.choice([15, 31, 63])),
                   uniquenessRatio=int(rng.choice([0, 10])), speckleWindowSize=int(rng.choice([0, 50])),
                   speckleRange=2, mode=mode)
        l = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
        r = np.roll(l, -int(rng.integers(1, 9)), axis=1)
        r = np.clip(r.astype(np.int16) + rng.integers(-10, 11, r.shape), 0, 255).astype(np.uint8)
        got = svb.StereoSGBM_create(**prm).compute(l, r)
        ref = oracle.sgbm_compute(l, r, **prm)
        assert np.array_equal(got, ref), (H, W, prm, int((got != ref).sum()))


def test_sgbm_3way_ties_and_min_disparity_vs_oracle():
    """SGBM_3WAY where cv2 has its own rules (pinned in the oracle by live cv2, 270/270 random cases): cost ties
    without a uniqueness test (8-lane argmin), minDisparity on both sides of zero.  Coarse grey levels make ties
    frequent."""
    _torch()
    rng = np.random.default_rng(2024)
    for it in range(40):
        D = int(rng.choice([16, 32, 48, 64, 96, 160]))
        bs = int(rng.choice([3, 5, 7]))
        minD = int(rng.integers(-16, 17))
        H = int(rng.integers(30, 80))
        W = int(rng.integers(D + abs(minD) + bs + 8, D + abs(minD) + 100))
        prm = dict(minDisparity=minD, numDisparities=D, blockSize=bs, P1=8 * bs * bs, P2=32 * bs * bs, disp12MaxDiff=1,
                   preFilterCap=63, uniquenessRatio=int(rng.choice([0, 0, 10])), speckleWindowSize=int(rng.choice([0, 30])),
                   speckleRange=2, mode=2)
        l, r, _ = make_pair(H, W, max(D, 16), seed=3000 + it, noise=int(rng.choice([0, 6])))
        if rng.random() < 0.5:
            l, r = (l // 32 * 32).astype(np.uint8), (r // 32 * 32).astype(np.uint8)
        got = svb.StereoSGBM_create(**prm).compute(l, r)
        ref = oracle.sgbm_compute(l, r, **prm)
        assert np.array_equal(got, ref), (H, W, prm, int((got != ref).sum()))


@pytest.mark.parametrize("D,bs,mode", [(256, 3, 1), (256, 7, 0), (256, 9, 1), (320, 7, 2), (384, 3, 0), (512, 7, 1), (208, 9, 0),
                                       (64, 7, 0), (48, 9, 2), (128, 7, 1), (256, 11, 0)])
def test_sgbm_block_sizes_at_wide_vectors_vs_oracle(D, bs, mode):
    """The cost kernels have instances with blockSize 3 / 5 / 7 / 9 built in (unrolled box sums) and a generic one
    (blockSize 11 here): every vector width they are dispatched at, including a padded one (208)."""
    _torch()
    H, W = 40, D + 90
    l, r, _ = make_pair(H, W, D, seed=40 + bs)
    prm = dict(minDisparity=0, numDisparities=D, blockSize=bs, P1=8 * bs * bs, P2=32 * bs * bs, disp12MaxDiff=1,
               preFilterCap=63, uniquenessRatio=10, speckleWindowSize=50, speckleRange=2, mode=mode)
    got = svb.StereoSGBM_create(**prm).compute(np.stack([l, l]), np.stack([r, r]))
    ref = oracle.sgbm_compute(l, r, **prm)
    assert np.array_equal(got[0], ref) and np.array_equal(got[1], ref)
