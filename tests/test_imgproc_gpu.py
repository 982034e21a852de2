"""GPU: remap / gray / median / speckle / depth / reproject kernels against the oracle and
the cv2-made golden vectors."""
import numpy as np
import pytest

import oracle
import stereovision_b200 as svb
from stereovision_b200.synth import make_pair, to_bgr
from tests.golden_cases import SGBM_CASES

pytestmark = pytest.mark.gpu


def test_remap_and_gray_bit_exact(golden):
    l, _, _ = make_pair(120, 160, 16, seed=7)
    bgr = to_bgr(l)
    m1, m2 = golden["remap_map1"], golden["remap_map2"]
    assert np.array_equal(svb.remap(bgr, m1, m2, svb.INTER_LINEAR), golden["remap_bgr"])
    assert np.array_equal(svb.remap(l, m1, m2, svb.INTER_LINEAR), golden["remap_gray"])
    assert np.array_equal(svb.cvtColor(bgr, svb.COLOR_BGR2GRAY), golden["gray"])
    fused = svb.remap(bgr, m1, m2, svb.INTER_LINEAR, to_gray=True)
    assert np.array_equal(fused, oracle.bgr2gray(golden["remap_bgr"]))
    # batched, and maps that point outside the source (BORDER_CONSTANT 0)
    rng = np.random.default_rng(3)
    src = rng.integers(0, 256, (3, 50, 70, 3), dtype=np.uint8)
    mm1 = rng.integers(-5, 75, (40, 60, 2)).astype(np.int16)
    mm2 = rng.integers(0, 1024, (40, 60)).astype(np.uint16)
    out = svb.remap(src, mm1, mm2, svb.INTER_LINEAR)
    for b in range(3):
        assert np.array_equal(out[b], oracle.remap(src[b], mm1, mm2))
    # widths that are not a multiple of 4 take the one-pixel-per-thread kernel
    mm1 = rng.integers(-5, 75, (33, 57, 2)).astype(np.int16)
    mm2 = rng.integers(0, 1024, (33, 57)).astype(np.uint16)
    for im in (src, src[..., 1].copy()):
        out = svb.remap(im, mm1, mm2, svb.INTER_LINEAR)
        for b in range(3):
            assert np.array_equal(out[b], oracle.remap(im[b], mm1, mm2))
    g = svb.remap(src, mm1, mm2, svb.INTER_LINEAR, to_gray=True)
    assert np.array_equal(g[1], oracle.bgr2gray(oracle.remap(src[1], mm1, mm2)))


def test_remap_map_cache_follows_the_arrays():
    """The device copies of the maps are cached per array object: a second call reuses them, a new array (even at
    the same address) or an edited one is uploaded again."""
    from stereovision_b200 import cv_api
    rng = np.random.default_rng(4)
    src = rng.integers(0, 256, (50, 70), dtype=np.uint8)
    mm1 = rng.integers(0, 45, (40, 60, 2)).astype(np.int16)
    mm2 = rng.integers(0, 1024, (40, 60)).astype(np.uint16)
    a = svb.remap(src, mm1, mm2, svb.INTER_LINEAR)
    n = len(cv_api._map_cache)
    assert np.array_equal(svb.remap(src, mm1, mm2, svb.INTER_LINEAR), a) and len(cv_api._map_cache) == n
    mm1[...] = rng.integers(0, 45, mm1.shape)                 # edited in place: the sample changes
    assert np.array_equal(svb.remap(src, mm1, mm2, svb.INTER_LINEAR), oracle.remap(src, mm1, mm2))
    for _ in range(3):                                         # fresh arrays, possibly at a recycled address
        mm1 = rng.integers(0, 45, (40, 60, 2)).astype(np.int16)
        assert np.array_equal(svb.remap(src, mm1, mm2, svb.INTER_LINEAR), oracle.remap(src, mm1, mm2))
    del mm1, mm2
    import gc
    gc.collect()
    assert len(cv_api._map_cache) <= n                         # entries leave with their arrays


def test_median_and_speckle_bit_exact(golden):
    d16 = golden["med3_in"]
    assert np.array_equal(svb.medianBlur(d16, 3), golden["med3"])
    f = d16.astype(np.float32) / 16.0
    assert np.array_equal(svb.medianBlur(f, 5), golden["med5"])
    prm = SGBM_CASES[0]["prm"]
    sp = d16.copy()
    svb.filterSpeckles(sp, (prm["minDisparity"] - 1) * 16, 40, 16)
    assert np.array_equal(sp, golden["speckle"])
    rng = np.random.default_rng(9)
    for it in range(6):
        H, W = int(rng.integers(5, 90)), int(rng.integers(5, 130))
        img = (rng.integers(0, 6, (2, H, W)) * int(rng.choice([8, 40])) - 16).astype(np.int16)
        maxsz, maxdiff = int(rng.choice([3, 20, 200])), int(rng.choice([0, 8, 40]))
        got = svb.filterSpeckles(img.copy(), -16, maxsz, maxdiff)
        for b in range(2):
            assert np.array_equal(got[b], oracle.filter_speckles(img[b], -16, maxsz, maxdiff)), (H, W, maxsz, maxdiff)
        fm = rng.normal(0, 30, (2, H, W)).astype(np.float32)
        got5 = svb.medianBlur(fm, 5)
        got3 = svb.medianBlur(img, 3)
        for b in range(2):
            assert np.array_equal(got5[b], oracle.median5_f32(fm[b]))
            assert np.array_equal(got3[b], oracle.median3_s16(img[b]))


def test_reproject_within_1e4(golden):
    d16 = golden["med3_in"]
    Q = golden["Q2"]
    for got, ref in ((svb.reprojectImageTo3D(d16, Q), golden["reproj_s16"]),
                     (svb.reprojectImageTo3D(d16.astype(np.float32) / 16.0, Q, True), golden["reproj_f32"])):
        fin = np.isfinite(ref)
        assert np.array_equal(fin, np.isfinite(got))
        np.testing.assert_allclose(got[fin], ref[fin], rtol=1e-4, atol=1e-6)


def test_depth_tail_matches_numpy_restatement(golden):
    """depth_map.py:909-933 restated in NumPy float32 (the reference's own arithmetic)."""
    d16 = golden["med3_in"]
    depth, disparity = svb.disparity_to_depth(d16, fx=700, baseline=0.08, min_disp=0, min_depth=0.3, max_depth=2.0)
    disp_ref = oracle.median5_f32(d16.astype(np.float32) / 16.0)
    assert np.array_equal(disparity, disp_ref)
    with np.errstate(divide="ignore"):
        z = (700 * 0.08) / (disp_ref + 1e-6)
    zc = np.clip(z, 0.3, 2.0)
    valid = (disp_ref > 0) & (zc >= 0.3) & (zc <= 2.0)
    ref = np.where(valid, zc, 0)
    assert ref.dtype == np.float32
    np.testing.assert_allclose(depth, ref, rtol=1e-6, atol=0)


@pytest.mark.parametrize("tag", ["l5", "r5", "l8", "l12", "l0"])
def test_init_undistort_rectify_map_bit_exact(golden, tag):
    """cv2.initUndistortRectifyMap(..., CV_16SC2) on the device: 5-, 8-, 12-coefficient models,
    no distortion / no rotation; and the maps feed svb.remap without leaving the device."""
    from oracle.maps import init_undistort_rectify_map
    from tests.test_oracle_golden import map_case
    K, dist, R, P, size = map_case(golden, tag)
    m1, m2 = svb.initUndistortRectifyMap(K, dist, R, P, size, svb.CV_16SC2)
    assert m1.dtype == np.int16 and m1.shape == (size[1], size[0], 2) and m2.dtype == np.uint16
    assert np.array_equal(m1, golden[f"maps_{tag}_m1"]) and np.array_equal(m2, golden[f"maps_{tag}_m2"])
    o1, o2 = init_undistort_rectify_map(K, dist, R, P, size)
    assert np.array_equal(m1, o1) and np.array_equal(m2, o2)
    t1, t2 = svb.initUndistortRectifyMap(K, dist, R, P, size, as_tensor=True)
    l, _, _ = make_pair(size[1], size[0], 16, seed=9)
    assert np.array_equal(svb.remap(to_bgr(l), t1, t2, svb.INTER_LINEAR), oracle.remap(to_bgr(l), m1, m2))


def test_init_undistort_rectify_map_full_size_and_errors(golden):
    from oracle.maps import init_undistort_rectify_map
    K, dist, R, P, _ = None, None, None, None, None
    from tests.test_oracle_golden import map_case
    K, dist, R, P, _ = map_case(golden, "l5")
    s = 1920 / 200.0
    Kb = K.copy(); Kb[:2] *= s
    Pb = P.copy(); Pb[:2] *= s
    m1, m2 = svb.initUndistortRectifyMap(Kb, dist, R, Pb, (1920, 1344))
    o1, o2 = init_undistort_rectify_map(Kb, dist, R, Pb, (1920, 1344))
    assert np.array_equal(m1, o1) and np.array_equal(m2, o2)
    with pytest.raises(ValueError):
        svb.initUndistortRectifyMap(K, np.zeros(6), R, P, (64, 48))          # cv2 rejects 6 coefficients too
    with pytest.raises(NotImplementedError):
        svb.initUndistortRectifyMap(K, np.r_[np.zeros(12), 0.01, 0.0], R, P, (64, 48))


def test_resize_matches_golden_and_oracle(golden):
    """cv2.resize(INTER_LINEAR) on the device (fused_depth_map.py:473-476: both frames every processed frame;
    :1214-1217: float32 mono map): u8 bit-exact against cv2's goldens and the oracle, float32 within 1e-6 relative;
    NumPy and tensor inputs, batched."""
    import torch
    from tests.golden_cases import RESIZE_CASES, resize_case_input
    for i, c in enumerate(RESIZE_CASES):
        src = resize_case_input(c)
        ref = golden[f"resize_{i}"]
        got = svb.resize(src, c["dsize"], interpolation=svb.INTER_LINEAR)
        assert got.shape == ref.shape and got.dtype == ref.dtype, i
        if c["ch"] == "f32":
            assert float(np.abs(got - ref).max()) <= 1e-6 * float(np.abs(ref).max()), i
            assert float(np.abs(got - oracle.resize(src, c["dsize"])).max()) <= 1e-6 * float(np.abs(ref).max()), i
        else:
            assert np.array_equal(got, ref), (i, int((got != ref).sum()))
            assert np.array_equal(got, oracle.resize(src, c["dsize"])), i
        batch = torch.from_numpy(np.stack([src, src[::-1].copy()])).cuda()
        out = svb.resize(batch, c["dsize"]).cpu().numpy()
        assert out.shape[0] == 2
        if c["ch"] == "f32":
            assert float(np.abs(out[0] - ref).max()) <= 1e-6 * float(np.abs(ref).max())
        else:
            assert np.array_equal(out[0], ref) and np.array_equal(out[1], oracle.resize(src[::-1].copy(), c["dsize"]))
    # fx / fy form and cv2's argument checks
    src = resize_case_input(RESIZE_CASES[0])
    assert svb.resize(src, None, fx=0.33, fy=0.33).shape == (round(480 * 0.33), round(640 * 0.33), 3)
    with pytest.raises(ValueError):
        svb.resize(src, (0, 0))
    with pytest.raises(NotImplementedError):
        svb.resize(src, (10, 10), interpolation=3)


def test_resize_random_sizes_vs_oracle():
    rng = np.random.default_rng(11)
    for it in range(60):
        sh, sw = int(rng.integers(1, 200)), int(rng.integers(1, 260))
        if it % 4 == 0:
            dsize = (max(1, int(sw * 0.33)), max(1, int(sh * 0.33)))
        else:
            dsize = (int(rng.integers(1, 300)), int(rng.integers(1, 300)))
        ch = int(rng.choice([1, 3]))
        src = rng.integers(0, 256, (sh, sw) if ch == 1 else (sh, sw, 3), dtype=np.uint8)
        got = svb.resize(src, dsize)
        assert np.array_equal(got, oracle.resize(src, dsize)), (sh, sw, dsize, ch)
    # 1080p frame at the reference's scale
    src = rng.integers(0, 256, (1080, 1920, 3), dtype=np.uint8)
    dsize = (int(1920 * 0.33), int(1080 * 0.33))
    assert np.array_equal(svb.resize(src, dsize), oracle.resize(src, dsize))
