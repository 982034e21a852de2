"""Generates tests/golden/golden_display.npz and oracle/colormaps.npz.
  * depth_map.create_depth_map is IMPORTED from /root/reference and run on the seeded pair of
    tests/golden_cases.DROPIN_CASE, once with fx = 700 (float32 arithmetic) and once with a float64 camera matrix
    (float64 arithmetic): its third output is the TURBO depth colour map;
  * the display expressions that live inside the reference's main loop (depth_map.py:1187-1195, 1207-1209;
    fused_depth_map.py:3026-3028) cannot be imported, so they are evaluated here with the same NumPy / cv2 calls
    on the reference's outputs;
  * fused_depth_map.draw_depth_legend_meters is imported and run with cv2.putText recorded, which yields the
    legend's formatted depth range.
Runs in the build container only:  python tests/golden/make_golden_display.py
"""
import contextlib
import io
import os
import sys

import cv2
import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
sys.path.insert(0, ROOT)
from stereovision_b200.synth import make_pair, make_aux_depths, to_bgr  # noqa: E402
from tests.golden_cases import DISPLAY_CASE, dense_disparity  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    sys.path.insert(0, "/root/reference")
    with contextlib.redirect_stdout(io.StringIO()):
        import depth_map as ref
        import fused_depth_map as fref
    c = DISPLAY_CASE
    l, r, gt = make_pair(c["H"], c["W"], 64, seed=c["seed"])
    lb, rb = to_bgr(l), to_bgr(r)
    K = np.array([[c["fx"], 0, c["W"] / 2], [0, c["fx"], c["H"] / 2], [0, 0, 1]], np.float64)
    out = {"cv2_version": np.array(cv2.__version__)}
    with contextlib.redirect_stdout(io.StringIO()):
        depth32, disp, cmap32 = ref.create_depth_map(lb, rb, None, min_depth=c["min_depth"], max_depth=c["max_depth"])
        depth64, disp2, cmap64 = ref.create_depth_map(lb, rb, {"calibration_data": {"camera_matrix": K}},
                                                      min_depth=c["min_depth"], max_depth=c["max_depth"])
    assert depth32.dtype == np.float32 and depth64.dtype == np.float64 and np.array_equal(disp, disp2)
    out.update(disparity=disp, depth32=depth32, depth64=depth64, depth_colormap32=cmap32, depth_colormap64=cmap64)
    # the same function on a dense disparity field (every multiple of 1/16 a matcher can produce): cv2.medianBlur
    # is replaced for these two calls only, so that everything after it (depth_map.py:915-937) is the reference's
    dense = dense_disparity(c)
    real_median = cv2.medianBlur
    cv2.medianBlur = lambda src, k: dense.copy()
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            d32, dd, c32 = ref.create_depth_map(lb, rb, None, min_depth=0.3, max_depth=2.0)
            d64, _, c64 = ref.create_depth_map(lb, rb, {"calibration_data": {"camera_matrix": K}}, min_depth=0.3, max_depth=2.0)
    finally:
        cv2.medianBlur = real_median
    assert np.array_equal(dd, dense) and d64.dtype == np.float64
    out.update(dense_depth32=d32, dense_depth64=d64, dense_colormap32=c32, dense_colormap64=c64)
    disp = dense                                     # the display expressions below run on the dense field too
    depth32, depth64 = d32, d64
    c = dict(c, min_depth=0.3, max_depth=2.0)
    # depth_map.py:1187-1195 (min_depth / max_depth are the loop's current range)
    for tag, depth in (("32", depth32), ("64", depth64)):
        min_depth, max_depth = c["min_depth"], c["max_depth"]
        depth_display = np.zeros_like(depth)
        valid_mask = depth > 0
        assert np.any(valid_mask)
        depth_display[valid_mask] = np.clip(depth[valid_mask], min_depth, max_depth)
        depth_display = ((depth_display - min_depth) / (max_depth - min_depth) * 255).astype(np.uint8)
        out["depth_display_index" + tag] = depth_display
        out["depth_display_colormap" + tag] = cv2.applyColorMap(depth_display, cv2.COLORMAP_TURBO)
    # depth_map.py:1207-1209
    disparity_normalized = cv2.normalize(disp, None, 0, 255, cv2.NORM_MINMAX)
    out["disparity_normalized_f32"] = disparity_normalized
    disparity_normalized = disparity_normalized.astype(np.uint8)
    out["disparity_index"] = disparity_normalized
    out["disparity_colormap"] = cv2.applyColorMap(disparity_normalized, cv2.COLORMAP_JET)
    # fused_depth_map.py:3026-3028 on a MiDaS-like map with values outside 0..255
    midas, flow = make_aux_depths(gt, c["seed"])
    wide = (midas.astype(np.float32) - np.float32(midas.mean())) * np.float32(1.7) + np.float32(120.0)
    out["clip_colormap"] = cv2.applyColorMap(np.clip(wide, 0, 255).astype(np.uint8), cv2.COLORMAP_JET)
    # the legend (fused_depth_map.py:1876-2110): record what it prints
    fused_u8 = np.where(flow > 0, np.clip(flow, 90, 255), 0).astype(np.uint8)
    texts = []
    real_put = cv2.putText

    def spy(img, text, *a, **k):
        texts.append(text)
        return real_put(img, text, *a, **k)
    cv2.putText = spy
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            fref.draw_depth_legend_meters(np.zeros((c["H"], c["W"], 3), np.uint8),
                                          {"focal_length": c["fx"], "baseline": 6.0}, fused_u8)
    finally:
        cv2.putText = real_put
    metres = [t for t in texts if t.endswith(" м")]
    assert len(metres) == 2, texts
    out["legend_max_text"], out["legend_min_text"] = np.array(metres[0]), np.array(metres[1])
    out["legend_consts"] = np.array([fref.MIN_DISP_BASE, fref.NUM_DISP_BASE], np.float64)
    path = os.path.join(HERE, "golden_display.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes", metres)
    ramp = np.arange(256, dtype=np.uint8).reshape(1, 256)
    np.savez(os.path.join(ROOT, "oracle", "colormaps.npz"), jet=cv2.applyColorMap(ramp, cv2.COLORMAP_JET)[0],
             turbo=cv2.applyColorMap(ramp, cv2.COLORMAP_TURBO)[0])


if __name__ == "__main__":
    main()
