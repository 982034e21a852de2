"""CPU oracle for the display-side reductions -- TEST INFRASTRUCTURE ONLY.

NumPy restatement of the reference's own NumPy expressions
    depth_map.py:926-937      depth_clipped -> uint8 -> COLORMAP_TURBO
    depth_map.py:1187-1195    displayed depth
    depth_map.py:1207-1209    cv2.normalize(disparity, 0, 255, NORM_MINMAX).astype(uint8) -> COLORMAP_JET
    fused_depth_map.py:1997-2050   legend range
plus the two OpenCV calls in them: cv2.applyColorMap (a 256-entry BGR table; oracle/colormaps.npz holds the JET
and TURBO tables read out of cv2 4.13 by tests/golden/make_golden_display.py) and cv2.normalize NORM_MINMAX with
a CV_32F result (scale and shift rounded to float32, dst = fma(src, scale, shift)).  Pinned by
tests/golden/golden_display.npz.
"""
import os

import numpy as np

f32 = np.float32
COLORMAP_JET, COLORMAP_TURBO = 2, 20
_tables = None


def lut(colormap):
    global _tables
    if _tables is None:
        _tables = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "colormaps.npz"))
    return _tables[{COLORMAP_JET: "jet", COLORMAP_TURBO: "turbo"}[colormap]]


def apply_colormap(idx, colormap):
    return lut(colormap)[np.asarray(idx, np.uint8)]


def depth_index(disparity, fx=700, baseline=0.08, min_depth=0.3, max_depth=2.0):
    """depth_map.py:926-936; fx may be a NumPy float64 scalar, which promotes the expression (NEP 50)."""
    disparity = np.asarray(disparity, f32)
    depth = (fx * baseline) / (disparity + 1e-6)
    depth_clipped = np.clip(depth, min_depth, max_depth)
    return ((depth_clipped - min_depth) / (max_depth - min_depth) * 255).astype(np.uint8)


def depth_display_index(depth, min_depth=0.3, max_depth=2.0):
    """depth_map.py:1187-1192."""
    depth_display = np.zeros_like(depth)
    valid = depth > 0
    depth_display[valid] = np.clip(depth[valid], min_depth, max_depth)
    return ((depth_display - min_depth) / (max_depth - min_depth) * 255).astype(np.uint8)


def normalize_minmax_f32(src, lo=0.0, hi=255.0):
    """cv2.normalize(src f32, None, lo, hi, cv2.NORM_MINMAX) -> f32."""
    src = np.asarray(src, f32)
    smin, smax = np.float64(src.min()), np.float64(src.max())
    scale = f32((hi - lo) * (1.0 / (smax - smin) if smax - smin > np.finfo(np.float64).eps else 0.0))
    shift = f32(lo) - f32(smin * np.float64(scale))
    # fma(src, scale, shift): the double product of two float32 is exact, one rounding of the sum
    return (src.astype(np.float64) * np.float64(scale) + np.float64(shift)).astype(f32)


def disparity_index(disparity):
    """depth_map.py:1207-1208."""
    return normalize_minmax_f32(disparity, 0.0, 255.0).astype(np.uint8)


def legend_range(fused_u8):
    v = np.asarray(fused_u8)
    v = v[v > 0]
    return (int(v.min()), int(v.max())) if v.size else (0, 0)


def legend_depth_range(fused_u8, min_disparity, num_disparities, focal_length, baseline_m, calibration_scale=3.75):
    """fused_depth_map.py:1997-2050."""
    lo, hi = legend_range(fused_u8)
    norm_min, norm_max = (max(1, lo), max(1, hi)) if hi > 0 else (1, 255)
    disp_min = max(0.5, min_disparity + (norm_min / 255.0) * num_disparities)
    disp_max = max(0.5, min_disparity + (norm_max / 255.0) * num_disparities)
    return (max(((focal_length * baseline_m) / disp_max) * calibration_scale, 0.05),
            min(((focal_length * baseline_m) / disp_min) * calibration_scale, 30.0))
