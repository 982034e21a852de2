"""GPU: the rows-per-warp sweep kernels (k_wave: first sweep at numDisparities <= 128; k_hlast: horizontal-only last
sweep) at the shapes that stress their grouping: heights around multiples of the rows per warp (4 / 2) and of the rows per
band (60 / 36), a single row, a batch whose images end in a partial band, 3WAY's overlapping stripes, negative and
positive minDisparity.  Bit-exact against the oracle (oracle/sgbm_oracle.c, pinned by the cv2 goldens)."""
import numpy as np
import pytest

import oracle
import stereovision_b200 as svb
from stereovision_b200.synth import make_pair

pytestmark = pytest.mark.gpu


def _torch():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch


def _prm(D, mode, minD=0):
    return dict(minDisparity=minD, numDisparities=D, blockSize=5, P1=200, P2=800, disp12MaxDiff=1, preFilterCap=63,
                uniquenessRatio=10, speckleWindowSize=60, speckleRange=32, mode=mode)


@pytest.mark.parametrize("D", [64, 128])
@pytest.mark.parametrize("mode", [0, 1, 2, 3])
def test_heights_around_the_group_and_band_sizes(D, mode):
    _torch()
    rows_band = 60 if D == 64 else 36
    heights = [1, 2, 3, 5, rows_band - 1, rows_band, rows_band + 1, 2 * rows_band + 3]
    W = D + 40
    big_l, big_r, _ = make_pair(max(heights), W, D, seed=900 + D + mode)
    for k, H in enumerate(heights):
        l, r = np.ascontiguousarray(big_l[k:k + H]), np.ascontiguousarray(big_r[k:k + H])   # (crops: make_pair needs room for its scene)
        prm = _prm(D, mode)
        want = oracle.sgbm_compute(l, r, **prm)
        for force in (1, 0):                                  # both first-sweep kernels (a single image alone would pick k_sweep)
            m = svb.StereoSGBM_create(**prm)
            m.setDebugForceWave(force)
            got = m.compute(l, r)
            assert np.array_equal(got, want), f"D={D} mode={mode} H={H} wave={force}: {int((got != want).sum())} pixels differ"


@pytest.mark.parametrize("D,minD", [(64, -8), (64, 16), (128, -24), (128, 8)])
def test_min_disparity_and_batches_with_partial_bands(D, minD):
    torch = _torch()
    H, W, B = 83, D + abs(minD) + 57, 5                        # 83 = 60 + 23 = 2 * 36 + 11: every image ends in a partial band
    prs = [make_pair(H, W, D, seed=950 + b) for b in range(B)]
    L, R = np.stack([p[0] for p in prs]), np.stack([p[1] for p in prs])
    for mode in (0, 2):
        prm = _prm(D, mode, minD)
        m = svb.StereoSGBM_create(**prm)
        m.setDebugForceWave(1)
        got = m.compute(torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()).cpu().numpy()
        for b in (0, B - 1):
            want = oracle.sgbm_compute(L[b], R[b], **prm)
            assert np.array_equal(got[b], want), f"D={D} minD={minD} mode={mode} image {b}: {int((got[b] != want).sum())} pixels differ"
        assert np.array_equal(got[2], svb.StereoSGBM_create(**prm).compute(L[2], R[2]))


@pytest.mark.parametrize("D", [16, 48, 80, 112])
def test_padded_disparity_vectors(D):
    """numDisparities that do not fill the 64-wide vector blocks (D < Dp): pad disparities carry the wall through both
    sweep kernels of the first sweep."""
    _torch()
    H, W = 67, D + 70
    l, r, _ = make_pair(H, W, max(D, 32), seed=990 + D)
    for mode in (0, 1, 2):
        prm = _prm(D, mode)
        want = oracle.sgbm_compute(l, r, **prm)
        for force in (1, 0):
            m = svb.StereoSGBM_create(**prm)
            m.setDebugForceWave(force)
            got = m.compute(l, r)
            assert np.array_equal(got, want), f"D={D} mode={mode} wave={force}: {int((got != want).sum())} pixels differ"
