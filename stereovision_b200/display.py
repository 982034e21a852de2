"""Display-side reductions on libsvb.so (csrc/display.cu): the normalisations and colour-map
lookups that turn a depth / disparity map into the BGR image the reference shows with cv2.imshow or saves with
cv2.imwrite.  There is no window system here: the result is a device (or NumPy) buffer.

    applyColorMap              cv2.applyColorMap(u8, COLORMAP_JET | COLORMAP_TURBO)      fused_depth_map.py:1016, 1700
    clip_colormap              applyColorMap(np.clip(x, 0, 255).astype(u8), JET)        fused_depth_map.py:3026-3028
    depth_colormap             create_depth_map's third output                           depth_map.py:926-937
    depth_display_colormap     the displayed depth (zeros where depth <= 0)             depth_map.py:1187-1195
    disparity_colormap         normalize(NORM_MINMAX, 0..255).astype(u8) -> JET         depth_map.py:1207-1209
    legend_range               min / max over the non-zero pixels                       fused_depth_map.py:1997-2004
    legend_depth_range         ... turned into metres                                    fused_depth_map.py:2005-2050
"""
import ctypes

import numpy as np

from . import _lib
from .cv_api import _ensure_device, _restores_device, _ret, _stream_ptr, _to_dev, _torch
from .fusion import _scratch

COLORMAP_JET, COLORMAP_TURBO = 2, 20
_U8, _CLIP_F32, _DISP_F32, _DISP_F64, _DEPTH_F32, _DEPTH_F64, _MINMAX_F32 = range(7)


def _is_f64_scalar(x):
    """NEP 50: a NumPy float64 scalar (fx read from a camera matrix) promotes float32 arrays, a Python number does not."""
    return isinstance(x, np.floating) and not isinstance(x, (np.float32, np.float16))


@_restores_device
def _run(src, kind, colormap, params=None, want_index=False, batched=None):
    torch = _torch()
    s, was = _to_dev(src)
    want = {_U8: torch.uint8, _DEPTH_F64: torch.float64}.get(kind, torch.float32)
    if s.dtype != want:
        raise ValueError(f"expected {want}, got {s.dtype}")
    if s.ndim not in (2, 3):
        raise ValueError("expected [H,W] or [B,H,W]")
    B = 1 if s.ndim == 2 else int(s.shape[0])
    npix = int(s.shape[-2]) * int(s.shape[-1])
    _ensure_device(s.device.index)
    L = _lib.lib()
    bgr = torch.empty(tuple(s.shape) + (3,), dtype=torch.uint8, device=s.device)
    idx = torch.empty(s.shape, dtype=torch.uint8, device=s.device) if want_index else None
    p = None if params is None else (ctypes.c_double * 5)(*[float(v) for v in params])
    nb = L.svb_display_scratch_bytes(B)
    scr = _scratch(nb, s.device)
    _lib.check(L.svb_colormap_dev(s.data_ptr(), kind, B, npix, int(colormap), p, idx.data_ptr() if want_index else None,
                                  bgr.data_ptr(), scr.data_ptr(), nb, _stream_ptr()))
    if want_index:
        return _ret(bgr, was), _ret(idx, was)
    return _ret(bgr, was)


def applyColorMap(src, colormap):
    """cv2.applyColorMap(src u8, colormap) for COLORMAP_JET and COLORMAP_TURBO -> BGR u8 [...,3]."""
    return _run(src, _U8, colormap)


def clip_colormap(src, colormap=COLORMAP_JET):
    """cv2.applyColorMap(np.clip(src, 0, 255).astype(np.uint8), colormap) for float32 maps."""
    return _run(src, _CLIP_F32, colormap)


def depth_colormap(disparity, fx=700, baseline=0.08, min_depth=0.3, max_depth=2.0, colormap=COLORMAP_TURBO,
                   return_index=False):
    """depth_map.create_depth_map's depth_colormap from its float32 disparity (depth_map.py:926-937).  A NumPy
    float64 fx selects float64 arithmetic, as it does in the reference."""
    kind = _DISP_F64 if _is_f64_scalar(fx) else _DISP_F32
    return _run(disparity, kind, colormap, (float(fx) * baseline, min_depth, max_depth, 0, 0), return_index)


def depth_display_colormap(depth, min_depth=0.3, max_depth=2.0, colormap=COLORMAP_TURBO, return_index=False):
    """depth_map.py:1187-1195 for a float32 or float64 depth map with at least one depth > 0."""
    d, _ = _to_dev(depth)
    kind = _DEPTH_F64 if d.dtype == _torch().float64 else _DEPTH_F32
    return _run(depth, kind, colormap, (0, min_depth, max_depth, 0, 0), return_index)


def disparity_colormap(disparity, colormap=COLORMAP_JET, alpha=0, beta=255, return_index=False):
    """depth_map.py:1207-1209: cv2.normalize(disparity, None, 0, 255, NORM_MINMAX).astype(u8) -> applyColorMap."""
    return _run(disparity, _MINMAX_F32, colormap, (0, 0, 0, alpha, beta), return_index)


@_restores_device
def legend_range(fused_u8):
    """(min, max) over the non-zero pixels of a u8 map, (0, 0) if there are none -- int32 [2] or [B,2]."""
    torch = _torch()
    s, was = _to_dev(fused_u8)
    if s.dtype != torch.uint8 or s.ndim not in (2, 3):
        raise ValueError("expected uint8 [H,W] or [B,H,W]")
    B = 1 if s.ndim == 2 else int(s.shape[0])
    _ensure_device(s.device.index)
    L = _lib.lib()
    out = torch.empty((B, 2), dtype=torch.int32, device=s.device)
    nb = L.svb_display_scratch_bytes(B)
    scr = _scratch(nb, s.device)
    _lib.check(L.svb_range_nonzero_u8_dev(s.data_ptr(), B, int(s.shape[-2]) * int(s.shape[-1]), out.data_ptr(),
                                          scr.data_ptr(), nb, _stream_ptr()))
    if s.ndim == 2:
        out = out[0]
    return _ret(out, was)


def legend_depth_range(fused_u8, min_disparity, num_disparities, focal_length, baseline_m, calibration_scale=3.75):
    """(min_depth_m, max_depth_m) of the legend, fused_depth_map.py:1997-2050: the reduction runs on the device,
    the dozen scalar operations after it are host arithmetic like the reference's."""
    r = legend_range(fused_u8)
    r = r.cpu().numpy() if not isinstance(r, np.ndarray) else r
    if r.ndim != 1:
        raise ValueError("legend_depth_range takes one map")
    norm_min, norm_max = (max(1, int(r[0])), max(1, int(r[1]))) if r[1] > 0 else (1, 255)
    disp_min = max(0.5, min_disparity + (norm_min / 255.0) * num_disparities)
    disp_max = max(0.5, min_disparity + (norm_max / 255.0) * num_disparities)
    max_depth_m = min(((focal_length * baseline_m) / disp_min) * calibration_scale, 30.0)
    min_depth_m = max(((focal_length * baseline_m) / disp_max) * calibration_scale, 0.05)
    return min_depth_m, max_depth_m
