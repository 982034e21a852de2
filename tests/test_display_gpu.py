"""GPU: SURVEY 8(f) N4 (csrc/display.cu) through stereovision_b200.display, bit-exact against the golden
vectors (reference's create_depth_map colour maps, cv2's normalize / applyColorMap) and oracle/display.py."""
import os

import numpy as np
import pytest

from oracle import display as D
import stereovision_b200 as svb
from tests.conftest import ROOT
from tests.golden_cases import DISPLAY_CASE as C, dense_disparity, display_case_inputs

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gd():
    return np.load(os.path.join(ROOT, "tests", "golden", "golden_display.npz"))


def test_depth_colormap_matches_create_depth_map(gd):
    fx64 = np.float64(C["fx"])
    kw = dict(min_depth=C["min_depth"], max_depth=C["max_depth"])
    assert np.array_equal(svb.depth_colormap(gd["disparity"], 700, 0.08, **kw), gd["depth_colormap32"])
    assert np.array_equal(svb.depth_colormap(gd["disparity"], fx64, 0.08, **kw), gd["depth_colormap64"])
    dense = dense_disparity()
    bgr, idx = svb.depth_colormap(dense, 700, return_index=True)
    assert np.array_equal(bgr, gd["dense_colormap32"]) and np.array_equal(idx, D.depth_index(dense, 700))
    assert np.array_equal(svb.depth_colormap(dense, fx64), gd["dense_colormap64"])


def test_disparity_to_depth_float64_fx(gd):
    """create_depth_map with a float64 camera matrix returns float64 depth (depth_map.py:917, 926)."""
    d16 = np.round(dense_disparity() * 16).astype(np.int16)
    K = np.array([[C["fx"], 0, 0], [0, C["fx"], 0], [0, 0, 1.0]])
    depth, disp = svb.disparity_to_depth(d16, fx=K[0, 0], baseline=0.08, min_disp=0, min_depth=0.3, max_depth=2.0)
    assert depth.dtype == np.float64 and disp.dtype == np.float32
    from oracle import median5_f32
    dm = median5_f32(d16.astype(np.float32) / 16.0)
    assert np.array_equal(disp, dm)
    z = np.clip((K[0, 0] * 0.08) / (dm + 1e-6), 0.3, 2.0)
    assert np.array_equal(depth, np.where(dm > 0, z, 0))                  # float64 arithmetic: bit-equal
    depth32, _ = svb.disparity_to_depth(d16, fx=700, baseline=0.08, min_disp=0, min_depth=0.3, max_depth=2.0)
    assert depth32.dtype == np.float32


def test_display_expressions_match_cv2(gd):
    dense = dense_disparity()
    bgr, idx = svb.disparity_colormap(dense, return_index=True)
    assert np.array_equal(idx, gd["disparity_index"]) and np.array_equal(bgr, gd["disparity_colormap"])
    for tag in ("32", "64"):
        bgr, idx = svb.depth_display_colormap(gd["dense_depth" + tag], 0.3, 2.0, return_index=True)
        assert np.array_equal(idx, gd["depth_display_index" + tag])
        assert np.array_equal(bgr, gd["depth_display_colormap" + tag])
    wide, fused = display_case_inputs()
    assert np.array_equal(svb.clip_colormap(wide), gd["clip_colormap"])
    assert np.array_equal(svb.applyColorMap(fused, svb.COLORMAP_JET), D.apply_colormap(fused, D.COLORMAP_JET))
    assert np.array_equal(svb.applyColorMap(fused, svb.COLORMAP_TURBO), D.apply_colormap(fused, D.COLORMAP_TURBO))
    with pytest.raises(NotImplementedError):
        svb.applyColorMap(fused, 4)


def test_batches_ragged_sizes_and_flat_images():
    rng = np.random.default_rng(12)
    for shape in [(1, 1), (3, 5), (2, 7, 9), (3, 16, 12), (1, 33, 31)]:
        x = rng.uniform(-50, 400, shape).astype(np.float32)
        bgr, idx = svb.disparity_colormap(x, return_index=True)
        imgs = x if x.ndim == 3 else x[None]
        want = np.stack([D.disparity_index(i) for i in imgs]).reshape(shape)
        assert np.array_equal(idx, want), shape
        assert np.array_equal(bgr, D.apply_colormap(want, D.COLORMAP_JET)), shape
        assert np.array_equal(svb.clip_colormap(x, svb.COLORMAP_TURBO),
                              D.apply_colormap(np.clip(x, 0, 255).astype(np.uint8), D.COLORMAP_TURBO))
    flat = np.full((6, 10), 3.5, np.float32)
    assert not svb.disparity_colormap(flat, return_index=True)[1].any()    # scale 0 -> every index 0


def test_legend_range(gd):
    _, fused = display_case_inputs()
    assert tuple(svb.legend_range(fused)) == D.legend_range(fused)
    mn, nd = gd["legend_consts"]
    lo, hi = svb.legend_depth_range(fused, mn, nd, C["fx"], 6.0 / 100.0)
    assert f"{hi:.2f} м" == str(gd["legend_max_text"]) and f"{lo:.2f} м" == str(gd["legend_min_text"])
    batch = np.stack([fused, np.zeros_like(fused), np.full_like(fused, 7)])
    assert np.array_equal(svb.legend_range(batch), np.array([D.legend_range(b) for b in batch], np.int32))
    assert svb.legend_depth_range(np.zeros((4, 5), np.uint8), 0, 320, 500, 0.06) == D.legend_depth_range(
        np.zeros((4, 5), np.uint8), 0, 320, 500, 0.06)
