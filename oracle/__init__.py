"""CPU oracle -- TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package; the product
(``stereovision_b200``) never does.  It wraps ``oracle/liboracle.so`` (plain C,
built by ``oracle/Makefile``) with ctypes and adds NumPy restatements of the
reference's float glue code.  Every function cites the reference file:line it
follows; the integer kernels follow SURVEY.md Appendix A (OpenCV 4.13.0 behaviour).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

MODE_SGBM, MODE_HH, MODE_SGBM_3WAY, MODE_HH4 = 0, 1, 2, 3


class _SgbmParams(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int) for n in (
        "minDisparity", "numDisparities", "blockSize", "P1", "P2", "disp12MaxDiff",
        "preFilterCap", "uniquenessRatio", "speckleWindowSize", "speckleRange", "mode")]


class _BmParams(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int) for n in (
        "minDisparity", "numDisparities", "blockSize", "preFilterCap", "textureThreshold",
        "uniquenessRatio", "speckleWindowSize", "speckleRange", "disp12MaxDiff", "preFilterType", "preFilterSize")]


def build(force=False):
    """Compile oracle/*.c -> oracle/liboracle.so (gcc only)."""
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith(".c")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build())
    return _LIB


def _p(a, t):
    return a.ctypes.data_as(ctypes.POINTER(t))


def sgbm_compute(left, right, minDisparity=0, numDisparities=16, blockSize=3, P1=0, P2=0,
                 disp12MaxDiff=0, preFilterCap=0, uniquenessRatio=0, speckleWindowSize=0,
                 speckleRange=0, mode=MODE_SGBM, stages=False):
    """cv2.StereoSGBM_create(...).compute(left, right) restated (depth_map.py:894-909).

    Returns int16 [H,W]; with stages=True also a dict with C (int16 [H,width1,D]),
    S (uint16) and the raw disparity before median/speckle."""
    left = np.ascontiguousarray(left, np.uint8)
    right = np.ascontiguousarray(right, np.uint8)
    assert left.shape == right.shape and (left.ndim == 2 or (left.ndim == 3 and left.shape[2] == 3))
    H, W = left.shape[:2]
    cn = 1 if left.ndim == 2 else 3
    prm = _SgbmParams(minDisparity, numDisparities, blockSize, P1, P2, disp12MaxDiff,
                      preFilterCap, uniquenessRatio, speckleWindowSize, speckleRange, mode)
    disp = np.empty((H, W), np.int16)
    maxD = minDisparity + numDisparities
    width1 = W + min(minDisparity, 0) - max(maxD, 0)
    C = S = None
    raw = np.empty((H, W), np.int16)
    if stages and width1 > 0:
        C = np.empty((H, width1, numDisparities), np.int16)
        S = np.empty((H, width1, numDisparities), np.uint16)
    rc = lib().orc_sgbm_compute_cn(
        _p(left, ctypes.c_uint8), _p(right, ctypes.c_uint8), H, W, cn, ctypes.byref(prm),
        _p(disp, ctypes.c_int16),
        _p(C, ctypes.c_int16) if C is not None else None,
        _p(S, ctypes.c_uint16) if S is not None else None,
        _p(raw, ctypes.c_int16))
    if rc != 0:
        raise ValueError(f"orc_sgbm_compute rejected the parameters (code {rc})")
    if stages:
        return disp, {"C": C, "S": S, "raw": raw}
    return disp


def remap(src, map1, map2):
    """cv2.remap(src, map1 CV_16SC2, map2 CV_16UC1, INTER_LINEAR) (depth_map.py:815-826)."""
    src = np.ascontiguousarray(src, np.uint8)
    C = 1 if src.ndim == 2 else src.shape[2]
    map1 = np.ascontiguousarray(map1, np.int16)
    map2 = np.ascontiguousarray(map2, np.uint16)
    Ho, Wo = map2.shape
    dst = np.empty((Ho, Wo) if src.ndim == 2 else (Ho, Wo, C), np.uint8)
    lib().orc_remap_u8(_p(src, ctypes.c_uint8), src.shape[0], src.shape[1], C,
                       _p(map1, ctypes.c_int16), _p(map2, ctypes.c_uint16), Ho, Wo,
                       _p(dst, ctypes.c_uint8))
    return dst


def resize(src, dsize):
    """cv2.resize(src, dsize=(W, H), interpolation=cv2.INTER_LINEAR): u8 with 1 or 3 channels (bit-exact,
    fused_depth_map.py:473-476, depth_map.py:809-811) or float32 single channel (fused_depth_map.py:1214-1217)."""
    Wd, Hd = int(dsize[0]), int(dsize[1])
    if src.dtype == np.float32:
        src = np.ascontiguousarray(src)
        assert src.ndim == 2
        dst = np.empty((Hd, Wd), np.float32)
        lib().orc_resize_f32(_p(src, ctypes.c_float), src.shape[0], src.shape[1], _p(dst, ctypes.c_float), Hd, Wd)
        return dst
    src = np.ascontiguousarray(src, np.uint8)
    C = 1 if src.ndim == 2 else src.shape[2]
    dst = np.empty((Hd, Wd) if src.ndim == 2 else (Hd, Wd, C), np.uint8)
    lib().orc_resize_u8(_p(src, ctypes.c_uint8), src.shape[0], src.shape[1], C, _p(dst, ctypes.c_uint8), Hd, Wd)
    return dst


def bgr2gray(bgr):
    """cv2.cvtColor(img, COLOR_BGR2GRAY) (depth_map.py:871-879)."""
    bgr = np.ascontiguousarray(bgr, np.uint8)
    out = np.empty(bgr.shape[:-1], np.uint8)
    lib().orc_bgr2gray(_p(bgr, ctypes.c_uint8), ctypes.c_size_t(out.size), _p(out, ctypes.c_uint8))
    return out


def median3_s16(a):
    a = np.ascontiguousarray(a, np.int16)
    out = np.empty_like(a)
    lib().orc_median3_s16(_p(a, ctypes.c_int16), a.shape[0], a.shape[1], _p(out, ctypes.c_int16))
    return out


def median5_f32(a):
    """cv2.medianBlur(disparity_f32, 5) (depth_map.py:912)."""
    a = np.ascontiguousarray(a, np.float32)
    out = np.empty_like(a)
    lib().orc_median5_f32(_p(a, ctypes.c_float), a.shape[0], a.shape[1], _p(out, ctypes.c_float))
    return out


def filter_speckles(img, newVal, maxSpeckleSize, maxDiff):
    """cv2.filterSpeckles on int16 (returns a filtered copy)."""
    out = np.ascontiguousarray(img, np.int16).copy()
    lib().orc_filter_speckles_s16(_p(out, ctypes.c_int16), out.shape[0], out.shape[1],
                                  int(newVal), int(maxSpeckleSize), int(maxDiff))
    return out


def reproject_image_to_3d(disp, Q, handleMissingValues=False):
    """cv2.reprojectImageTo3D(disp, Q) -- int16 input is NOT divided by 16."""
    d = np.ascontiguousarray(disp, np.float32)
    Q = np.ascontiguousarray(Q, np.float64)
    out = np.empty(d.shape + (3,), np.float32)
    lib().orc_reproject3d(_p(d, ctypes.c_float), d.shape[0], d.shape[1], _p(Q, ctypes.c_double),
                          int(bool(handleMissingValues)), _p(out, ctypes.c_float))
    return out


def bm_compute(left, right, numDisparities=64, blockSize=21, minDisparity=0, preFilterCap=31,
               textureThreshold=10, uniquenessRatio=15, speckleWindowSize=0, speckleRange=0, disp12MaxDiff=-1,
               preFilterType=1, preFilterSize=9):
    """cv2.StereoBM_create(numDisparities, blockSize).compute(left, right) restated."""
    left = np.ascontiguousarray(left, np.uint8)
    right = np.ascontiguousarray(right, np.uint8)
    assert left.shape == right.shape and left.ndim == 2
    H, W = left.shape
    prm = _BmParams(minDisparity, numDisparities, blockSize, preFilterCap, textureThreshold,
                    uniquenessRatio, speckleWindowSize, speckleRange, disp12MaxDiff, preFilterType, preFilterSize)
    disp = np.empty((H, W), np.int16)
    rc = lib().orc_bm_compute(_p(left, ctypes.c_uint8), _p(right, ctypes.c_uint8), H, W, ctypes.byref(prm),
                              _p(disp, ctypes.c_int16))
    if rc != 0:
        raise ValueError(f"orc_bm_compute rejected the parameters (code {rc})")
    return disp


def validate_disparity(disp, cost, minDisparity, numberOfDisparities, disp12MaxDiff=1):
    """cv2.validateDisparity restated (returns a validated copy)."""
    out = np.ascontiguousarray(disp, np.int16).copy()
    c = np.ascontiguousarray(cost, np.int32)
    assert out.ndim == 2 and c.shape == out.shape
    lib().orc_validate_disparity(_p(out, ctypes.c_int16), _p(c, ctypes.c_int32), out.shape[0], out.shape[1],
                                 int(minDisparity), int(numberOfDisparities), int(disp12MaxDiff))
    return out
