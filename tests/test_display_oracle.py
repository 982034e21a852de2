"""CPU: oracle/display.py against tests/golden/golden_display.npz -- the colour maps of the
reference's own create_depth_map, and the display expressions of its main loop evaluated with cv2."""
import os

import numpy as np
import pytest

from oracle import display as D
from tests.conftest import ROOT
from tests.golden_cases import DISPLAY_CASE as C, dense_disparity, display_case_inputs


@pytest.fixture(scope="module")
def gd():
    return np.load(os.path.join(ROOT, "tests", "golden", "golden_display.npz"))


def test_depth_colormap_matches_create_depth_map(gd):
    fx64 = np.float64(C["fx"])
    idx = D.depth_index(gd["disparity"], 700, 0.08, C["min_depth"], C["max_depth"])
    assert np.array_equal(D.apply_colormap(idx, D.COLORMAP_TURBO), gd["depth_colormap32"])
    idx = D.depth_index(gd["disparity"], fx64, 0.08, C["min_depth"], C["max_depth"])
    assert np.array_equal(D.apply_colormap(idx, D.COLORMAP_TURBO), gd["depth_colormap64"])
    dense = dense_disparity()
    assert np.array_equal(D.apply_colormap(D.depth_index(dense, 700), D.COLORMAP_TURBO), gd["dense_colormap32"])
    assert np.array_equal(D.apply_colormap(D.depth_index(dense, fx64), D.COLORMAP_TURBO), gd["dense_colormap64"])
    assert len(np.unique(gd["dense_colormap32"].reshape(-1, 3), axis=0)) == 256      # every table entry is exercised


def test_display_expressions_match_cv2(gd):
    dense = dense_disparity()
    assert np.array_equal(D.normalize_minmax_f32(dense), gd["disparity_normalized_f32"])
    assert np.array_equal(D.disparity_index(dense), gd["disparity_index"])
    assert np.array_equal(D.apply_colormap(D.disparity_index(dense), D.COLORMAP_JET), gd["disparity_colormap"])
    for tag in ("32", "64"):
        idx = D.depth_display_index(gd["dense_depth" + tag])
        assert np.array_equal(idx, gd["depth_display_index" + tag])                    # incl. the wrapped negatives
        assert np.array_equal(D.apply_colormap(idx, D.COLORMAP_TURBO), gd["depth_display_colormap" + tag])
    wide, _ = display_case_inputs()
    assert np.array_equal(D.apply_colormap(np.clip(wide, 0, 255).astype(np.uint8), D.COLORMAP_JET), gd["clip_colormap"])


def test_legend_range_matches_reference_text(gd):
    _, fused = display_case_inputs()
    mn, nd = gd["legend_consts"]
    lo, hi = D.legend_depth_range(fused, mn, nd, C["fx"], 6.0 / 100.0)
    assert f"{hi:.2f} м" == str(gd["legend_max_text"]) and f"{lo:.2f} м" == str(gd["legend_min_text"])
    assert D.legend_range(np.zeros((4, 5), np.uint8)) == (0, 0)
    assert D.legend_depth_range(np.zeros((4, 5), np.uint8), 0, 320, 500, 0.06) == (max(500 * 0.06 / 320 * 3.75, 0.05), 30.0)
