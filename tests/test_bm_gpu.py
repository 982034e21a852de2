"""GPU: StereoBM (BASELINE config 2) against the oracle and the cv2-made golden vectors."""
import numpy as np
import pytest

import oracle
import stereovision_b200 as svb
from stereovision_b200.synth import make_pair, make_batch
from tests.golden_cases import BM_CASES
from tests.test_oracle_golden import _bm_kwargs

pytestmark = pytest.mark.gpu


def _matcher(kw):
    m = svb.StereoBM_create(kw["numDisparities"], kw["blockSize"])
    for k, v in kw.items():
        if k not in ("numDisparities", "blockSize"):
            getattr(m, "set" + k[0].upper() + k[1:])(v)
    return m


@pytest.mark.parametrize("i", range(len(BM_CASES)))
def test_bm_matches_golden(golden, i):
    c = BM_CASES[i]
    l, r, _ = make_pair(c["H"], c["W"], max(c["numDisparities"], 16), seed=c["seed"], noise=c.get("noise", 6))
    got = _matcher(_bm_kwargs(c)).compute(l, r)
    assert got.dtype == np.int16
    assert np.array_equal(got, golden[f"bm_{i}"])


def test_bm_random_params_vs_oracle():
    rng = np.random.default_rng(77)
    for it in range(48):
        D = int(rng.choice([16, 32, 64, 96, 128]))
        bs = int(rng.choice([5, 7, 9, 15, 21]))
        minD = int(rng.choice([0, 0, -4, -20, 1, 5, 13]))     # > 0: cv2's row-end spill (oracle/bm_oracle.c)
        H = int(rng.integers(bs + 6, 80))
        W = int(rng.integers(D + abs(minD) + bs + 8, D + abs(minD) + 140))
        kw = dict(numDisparities=D, blockSize=bs, minDisparity=minD, preFilterCap=int(rng.choice([5, 31, 63])),
                  textureThreshold=int(rng.choice([0, 10, 200])), uniquenessRatio=int(rng.choice([0, 5, 15])),
                  speckleWindowSize=int(rng.choice([0, 30])), speckleRange=int(rng.choice([1, 4])),
                  disp12MaxDiff=int(rng.choice([-1, -1, 0, 1, 4])), preFilterType=int(rng.choice([1, 1, 0])),
                  preFilterSize=int(rng.choice([5, 9, 21])))
        if kw["disp12MaxDiff"] >= 0 and bs * bs * 2 * kw["preFilterCap"] > 32767:
            kw["preFilterCap"] = 31
        l, r, _ = make_pair(H, W, max(D, 16), seed=700 + it, noise=int(rng.choice([0, 6, 25])))
        got = _matcher(kw).compute(l, r)
        ref = oracle.bm_compute(l, r, **kw)
        assert np.array_equal(got, ref), (H, W, kw, int((got != ref).sum()))


def test_bm_config2_shape_batch_and_properties():
    """BASELINE config 2: 640x480, D=64, bs=15 -- batch == singles, one image vs the oracle."""
    import torch
    B, H, W, D = 6, 480, 640, 64
    L, R = make_batch(B, H, W, D, seed0=5)
    m = svb.StereoBM_create(D, 15)
    out = m.compute(L, R)
    assert np.array_equal(out[0], oracle.bm_compute(L[0], R[0], D, 15))
    dev = m.compute(torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()).cpu().numpy()
    assert np.array_equal(out, dev)
    assert np.array_equal(out[3], m.compute(L[3], R[3]))
    assert np.all(out[:, :7, :] == -16) and np.all(out[:, :, :D - 1 + 7] == -16)   # outside the valid ROI


def test_validate_disparity_matches_golden(golden):
    """cv2.validateDisparity as a stand-alone call: int16 and int32 cost maps, ties, NumPy in-place contract."""
    import torch
    d = golden["validate_in"]
    for k, (cost, d12) in enumerate(((golden["validate_cost16"], 1), (golden["validate_cost32"], 0),
                                     (golden["validate_cost16"], 4))):
        v = d.copy()
        ret = svb.validateDisparity(v, cost, 0, 16, d12)
        assert ret is v and np.array_equal(v, golden[f"validate_out_{k}"]), k
    dt = torch.from_numpy(np.stack([d, d[::-1].copy()])).cuda()
    ct = torch.from_numpy(np.stack([golden["validate_cost16"], golden["validate_cost16"][::-1].copy()])).cuda()
    out = svb.validateDisparity(dt, ct, 0, 16, 1).cpu().numpy()
    assert np.array_equal(out[0], golden["validate_out_0"]) and np.array_equal(out[1], golden["validate_out_0"][::-1])


def test_bm_errors_like_cv2():
    with pytest.raises(ValueError):
        svb.StereoBM_create(24, 15).compute(np.zeros((40, 100), np.uint8), np.zeros((40, 100), np.uint8))
    with pytest.raises(ValueError):
        svb.StereoBM_create(16, 8).compute(np.zeros((40, 100), np.uint8), np.zeros((40, 100), np.uint8))
    # disparity range wider than the image: cv2 returns an all-invalid map, no error
    out = svb.StereoBM_create(64, 5).compute(np.zeros((40, 50), np.uint8), np.zeros((40, 50), np.uint8))
    assert np.all(out == -16)
