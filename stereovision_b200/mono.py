"""The data-parallel code around the two mono estimators of fused_depth_map.py,
on libsvb.so (csrc/fuse.cu).  The networks / optical flow themselves are the caller's: these functions
take their raw outputs.  NumPy in -> NumPy out, CUDA tensors in -> CUDA tensors out; a leading batch
dimension is accepted.

    midas_postprocess      fused_depth_map.py:1134-1158  (bicubic upsample, 0..255 normalise, contrast confidence)
    contrast_confidence    fused_depth_map.py:1150-1156
    flow_residual_depth    fused_depth_map.py:1386-1403 with the expected flow of 1491-1497
    bilateralFilter        cv2.bilateralFilter on float32 maps (fused_depth_map.py:1400, 1409, 1690)
"""
import ctypes

import numpy as np

from . import _lib
from .cv_api import _ensure_device, _restores_device, _is_tensor, _ret, _stream_ptr, _to_dev, _torch, cvtColor
from .fusion import _f32_maps, _scratch


@_restores_device
def bilateralFilter(src, d, sigmaColor, sigmaSpace):
    """cv2.bilateralFilter(src, d, sigmaColor, sigmaSpace) for float32 [H,W] / [B,H,W]; d odd, <= 15
    (the reference only uses 9, 75, 75)."""
    torch = _torch()
    (s,), was, B, H, W = _f32_maps(src)
    _ensure_device(s.device.index)
    L = _lib.lib()
    d = int(d)
    if d <= 0:                                           # cv2: derived from sigmaSpace
        d = max(int(round(float(sigmaSpace) * 1.5)), 1) * 2 + 1
    out = torch.empty_like(s)
    nb = L.svb_bilateral_scratch_bytes(B)
    scr = _scratch(nb, s.device)
    _lib.check(L.svb_bilateral_f32_dev(s.data_ptr(), B, H, W, d, float(sigmaColor), float(sigmaSpace), out.data_ptr(),
                                       scr.data_ptr(), nb, _stream_ptr()))
    return _ret(out, was)


@_restores_device
def upsample_normalize(prediction, size):
    """F.interpolate(prediction[:, None], size=size, mode="bicubic", align_corners=False) followed by the
    per-image 0..255 min-max normalisation of fused_depth_map.py:1134-1145.  -> float32 [B,]H,W."""
    torch = _torch()
    (p,), was, B, Hi, Wi = _f32_maps(prediction)
    Ho, Wo = int(size[0]), int(size[1])
    _ensure_device(p.device.index)
    L = _lib.lib()
    out = torch.empty((B, Ho, Wo), dtype=torch.float32, device=p.device)
    nb = L.svb_midas_scratch_bytes(B)
    scr = _scratch(nb, p.device)
    _lib.check(L.svb_midas_upsample_normalize_dev(p.data_ptr(), B, Hi, Wi, out.data_ptr(), Ho, Wo, scr.data_ptr(), nb,
                                                  _stream_ptr()))
    if p.ndim == 2:
        out = out[0]
    return _ret(out, was)


@_restores_device
def contrast_confidence(frame):
    """cv2.normalize(cv2.absdiff(gray, cv2.GaussianBlur(gray, (5, 5), 0)), None, 0, 1, cv2.NORM_MINMAX)
    (fused_depth_map.py:1150-1156).  frame: u8 gray [H,W] / [B,H,W] or BGR [H,W,3] / [B,H,W,3]; -> u8."""
    torch = _torch()
    g, was = _to_dev(frame)
    if g.dtype != torch.uint8:
        raise ValueError("expected a uint8 frame")
    if g.shape[-1] == 3 and g.ndim in (3, 4):
        g = cvtColor(g)
    if g.ndim not in (2, 3):
        raise ValueError("expected [H,W], [B,H,W], [H,W,3] or [B,H,W,3]")
    B = 1 if g.ndim == 2 else int(g.shape[0])
    H, W = int(g.shape[-2]), int(g.shape[-1])
    _ensure_device(g.device.index)
    L = _lib.lib()
    out = torch.empty_like(g)
    nb = L.svb_midas_scratch_bytes(B)
    scr = _scratch(nb, g.device)
    _lib.check(L.svb_contrast_confidence_u8_dev(g.data_ptr(), B, H, W, out.data_ptr(), scr.data_ptr(), nb, _stream_ptr()))
    return _ret(out, was)


def midas_postprocess(prediction, frame):
    """Everything of MiDaSDepthEstimator.estimate_depth after the network (fused_depth_map.py:1134-1158):
    -> (depth_normalized f32 0..255 at the frame's size, confidence u8)."""
    f, _ = _to_dev(frame)
    H, W = (int(f.shape[-3]), int(f.shape[-2])) if f.shape[-1] == 3 and f.ndim >= 3 else (int(f.shape[-2]), int(f.shape[-1]))
    p, was = _to_dev(prediction, np.float32) if not _is_tensor(prediction) else _to_dev(prediction)
    depth = upsample_normalize(p, (H, W))
    conf = contrast_confidence(f)
    return _ret(depth, was), _ret(conf, _is_tensor(frame))


@_restores_device
def flow_residual_depth(flow, homography, stable_depth=None, camera_moving=True):
    """OpticalFlowDepthEstimator.estimate_depth after calcOpticalFlowFarneback / findHomography
    (fused_depth_map.py:1386-1403): residual of the dense flow against the ego-motion homography ->
    1 / (|residual| + 0.5), the exponential average and the 9/75/75 bilateral filter.
    flow: f32 [H,W,2] or [B,H,W,2]; homography: 3x3 (or [B,3,3]) float64.
    -> (depth_smoothed, stable_depth, current_depth); pass stable_depth back in on the next frame."""
    torch = _torch()
    fl, was = _to_dev(flow, np.float32) if not _is_tensor(flow) else _to_dev(flow)
    if fl.dtype != torch.float32 or fl.shape[-1] != 2 or fl.ndim not in (3, 4):
        raise ValueError("expected float32 flow [H,W,2] or [B,H,W,2]")
    batched = fl.ndim == 4
    B = int(fl.shape[0]) if batched else 1
    H, W = int(fl.shape[-3]), int(fl.shape[-2])
    hm = np.ascontiguousarray(np.broadcast_to(np.asarray(homography, np.float64).reshape(-1, 3, 3), (B, 3, 3)))
    _ensure_device(fl.device.index)
    L = _lib.lib()
    cur = torch.empty((B, H, W), dtype=torch.float32, device=fl.device)
    if stable_depth is None:
        stable, valid = torch.empty_like(cur), 0
    else:
        st, _ = _to_dev(stable_depth, np.float32) if not _is_tensor(stable_depth) else _to_dev(stable_depth)
        stable, valid = st.reshape(B, H, W).clone(), 1
    alpha = 0.9 if camera_moving else 0.99
    _lib.check(L.svb_flow_residual_depth_dev(fl.data_ptr(), B, H, W, hm.ctypes.data_as(ctypes.POINTER(ctypes.c_double)),
                                             cur.data_ptr(), stable.data_ptr(), valid, alpha, _stream_ptr()))
    smooth = bilateralFilter(cur, 9, 75, 75)
    if not batched:
        smooth, stable, cur = smooth[0], stable[0], cur[0]
    return _ret(smooth, was), _ret(stable, was), _ret(cur, was)
