"""Calibration-file path of the reference: pickle -> stereoRectify -> 2 x initUndistortRectifyMap -> the dict that
apply_stereo_rectification consumes.

    depth_map.load_stereo_calibration                    depth_map.py:591-668
    fused_depth_map.load_stereo_calibration_with_scaling fused_depth_map.py:307-442 (K scaled by scale_factor;
                                                         re-run on '+' / '-' at :3339, :3355)

The rectification maps (the per-pixel work) are computed on the device by svb.initUndistortRectifyMap
(k_init_undistort_rectify_map, float64, bit-equal to cv2).  cv2.stereoRectify itself is a few 3x3 products, two
Rodrigues conversions and 2 x (4 + 81) undistorted points: it runs on the host in NumPy float64 below (`stereoRectify`),
restating OpenCV 4.13 calib3d/src/calibration.cpp (cvStereoRectify, icvGetRectangles, cvUndistortPoints with its
fixed 5 iterations, point sets held in float32 as OpenCV holds them).  No cv2 import anywhere in the product.
"""
import os
import pickle

import numpy as np

CALIB_ZERO_DISPARITY = 1024
CALIBRATION_KEYS = ("mtx_left", "dist_left", "mtx_right", "dist_right", "R", "T", "img_size")   # depth_map.py:613-619


def Rodrigues(src):
    """cv2.Rodrigues without the Jacobian: 3-vector -> 3x3 matrix, or 3x3 matrix -> 3-vector (column)."""
    a = np.asarray(src, np.float64)
    if a.size == 3:
        r = a.reshape(3)
        theta = float(np.sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]))
        if theta < np.finfo(np.float64).eps:
            return np.eye(3)
        c, s = np.cos(theta), np.sin(theta)
        c1 = 1.0 - c
        x, y, z = r / theta
        rrt = np.array([[x * x, x * y, x * z], [x * y, y * y, y * z], [x * z, y * z, z * z]])
        rx = np.array([[0, -z, y], [z, 0, -x], [-y, x, 0]])
        return c * np.eye(3) + c1 * rrt + s * rx
    if a.shape != (3, 3):
        raise ValueError("Rodrigues: expected a 3-vector or a 3x3 matrix")
    U, _, Vt = np.linalg.svd(a)
    R = U @ Vt                                                # nearest rotation, as OpenCV orthogonalises first
    r = np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
    s = float(np.sqrt((r[0] * r[0] + r[1] * r[1] + r[2] * r[2]) * 0.25))
    c = (R[0, 0] + R[1, 1] + R[2, 2] - 1) * 0.5
    c = 1.0 if c > 1.0 else (-1.0 if c < -1.0 else c)
    theta = float(np.arccos(c))
    if s < 1e-5:
        if c > 0:
            return np.zeros((3, 1))
        t = (R[0, 0] + 1) * 0.5
        rx = np.sqrt(max(t, 0.0))
        t = (R[1, 1] + 1) * 0.5
        ry = np.sqrt(max(t, 0.0)) * (-1.0 if R[0, 1] < 0 else 1.0)
        t = (R[2, 2] + 1) * 0.5
        rz = np.sqrt(max(t, 0.0)) * (-1.0 if R[0, 2] < 0 else 1.0)
        if abs(rx) < abs(ry) and abs(rx) < abs(rz) and (R[1, 2] > 0) != (ry * rz > 0):
            rz = -rz
        v = np.array([rx, ry, rz])
        v *= theta / np.sqrt(v @ v)
        return v.reshape(3, 1)
    return (r * (theta / (2 * s))).reshape(3, 1)


def _dist12(dist):
    k = np.zeros(14)
    if dist is not None:
        d = np.asarray(dist, np.float64).ravel()
        if d.size not in (0, 4, 5, 8, 12, 14):
            raise ValueError("distortion coefficients: 4, 5, 8, 12 or 14 values")
        if d.size == 14 and (d[12] != 0 or d[13] != 0):
            raise NotImplementedError("the tilted sensor model is not on the reference's path")
        k[:d.size] = d
    return k


def undistort_points(pts, K, dist, R=None, P=None):
    """cv2.undistortPoints as stereoRectify uses it (C entry point: exactly 5 fixed-point iterations, no epsilon
    test).  pts: [n,2]; returns float64 [n,2] (callers round to float32 where OpenCV stores float32)."""
    K = np.asarray(K, np.float64)
    k = _dist12(dist)
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    ifx, ify = 1.0 / fx, 1.0 / fy
    RR = np.eye(3) if R is None else np.asarray(R, np.float64)
    if P is not None:
        RR = np.asarray(P, np.float64)[:3, :3] @ RR
    out = np.empty((len(pts), 2))
    for i, (u, v) in enumerate(np.asarray(pts, np.float64)):
        x0 = x = (u - cx) * ifx
        y0 = y = (v - cy) * ify
        if dist is not None:
            for _ in range(5):
                r2 = x * x + y * y
                icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2)
                if icdist < 0:
                    x, y = (u - cx) * ifx, (v - cy) * ify
                    break
                dX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2
                dY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2
                x = (x0 - dX) * icdist
                y = (y0 - dY) * icdist
        xx = RR[0, 0] * x + RR[0, 1] * y + RR[0, 2]
        yy = RR[1, 0] * x + RR[1, 1] * y + RR[1, 2]
        ww = 1.0 / (RR[2, 0] * x + RR[2, 1] * y + RR[2, 2])
        out[i] = (xx * ww, yy * ww)
    return out


def _f32(a):
    return np.asarray(a, np.float32).astype(np.float64)


def _rectangles(K, dist, R, P, size):
    """getRectangles: a 9 x 9 grid of image points through undistortPoints(R, P); inner = the largest axis-aligned
    rectangle inside the warped border, outer = its bounding box.  Double precision throughout, as OpenCV 4.13
    holds them (probed: a float32 grid is off by 3e-11 relative in the scale).  Returns (inner, outer) as (x, y, w, h)."""
    N = 9
    W, H = size
    pts = np.array([[x * (W - 1) / (N - 1), y * (H - 1) / (N - 1)] for y in range(N) for x in range(N)], np.float64)
    p = undistort_points(pts, K, dist, R, P).reshape(N, N, 2)
    ox0, ox1, oy0, oy1 = p[..., 0].min(), p[..., 0].max(), p[..., 1].min(), p[..., 1].max()
    ix0, ix1 = p[:, 0, 0].max(), p[:, N - 1, 0].min()
    iy0, iy1 = p[0, :, 1].max(), p[N - 1, :, 1].min()
    inner = tuple(float(v) for v in (ix0, iy0, ix1 - ix0, iy1 - iy0))
    outer = tuple(float(v) for v in (ox0, oy0, ox1 - ox0, oy1 - oy0))
    return inner, outer


def _rect_and(a, b):
    x0, y0 = max(a[0], b[0]), max(a[1], b[1])
    w, h = min(a[0] + a[2], b[0] + b[2]) - x0, min(a[1] + a[3], b[1] + b[3]) - y0
    return (x0, y0, w, h) if w > 0 and h > 0 else (0, 0, 0, 0)


def stereoRectify(cameraMatrix1, distCoeffs1, cameraMatrix2, distCoeffs2, imageSize, R, T, flags=CALIB_ZERO_DISPARITY,
                  alpha=-1, newImageSize=(0, 0)):
    """cv2.stereoRectify(K1, d1, K2, d2, (W, H), R, T, alpha=..., flags=...) -> (R1, R2, P1, P2, Q, roi1, roi2)
    (depth_map.py:625-630 with alpha=0 and the default flags; fused_depth_map.py:392-400 with
    CALIB_ZERO_DISPARITY spelled out).  Host float64; OpenCV 4.13 cvStereoRectify restated."""
    K = [np.asarray(cameraMatrix1, np.float64), np.asarray(cameraMatrix2, np.float64)]
    D = [distCoeffs1, distCoeffs2]
    nx, ny = int(imageSize[0]), int(imageSize[1])
    Rm = np.asarray(R, np.float64)
    Tv = np.asarray(T, np.float64).reshape(3)
    om = Rodrigues(Rm).reshape(3) if Rm.shape == (3, 3) else Rm.reshape(3).copy()
    r_r = Rodrigues(om * -0.5)                                 # rotate both cameras halfway
    t = r_r @ Tv
    idx = 0 if abs(t[0]) > abs(t[1]) else 1
    c, nt = t[idx], float(np.sqrt(t @ t))
    if not nt > 0:
        raise ValueError("stereoRectify: zero translation")
    uu = np.zeros(3)
    uu[idx] = 1.0 if c > 0 else -1.0
    ww = np.cross(t, uu)
    nw = float(np.sqrt(ww @ ww))
    if nw > 0:
        ww = ww * (np.arccos(abs(c) / nt) / nw)
    wR = Rodrigues(ww)
    R1 = wR @ r_r.T
    R2 = wR @ r_r
    t = R2 @ Tv
    new = (int(newImageSize[0]), int(newImageSize[1]))
    if new[0] * new[1] == 0:
        new = (nx, ny)
    ratio_x, ratio_y = new[0] / nx / 2, new[1] / ny / 2
    ratio = ratio_x if idx == 1 else ratio_y
    fc_new = (K[0][idx ^ 1, idx ^ 1] + K[1][idx ^ 1, idx ^ 1]) * ratio
    cc = [None, None]
    for k in range(2):
        corners = np.array([[(i % 2) * (nx - 1), (0 if i < 2 else 1) * (ny - 1)] for i in range(4)], np.float32)
        p = _f32(undistort_points(corners, K[k], D[k]))                    # float32 points, z = 1
        Rk = R1 if k == 0 else R2
        X = p[:, 0] * Rk[0, 0] + p[:, 1] * Rk[0, 1] + Rk[0, 2]
        Y = p[:, 0] * Rk[1, 0] + p[:, 1] * Rk[1, 1] + Rk[1, 2]
        Z = p[:, 0] * Rk[2, 0] + p[:, 1] * Rk[2, 1] + Rk[2, 2]
        iz = 1.0 / Z
        proj = _f32(np.stack([X * iz * fc_new, Y * iz * fc_new], 1))       # projectPoints into a float32 array
        cc[k] = [(nx - 1) / 2 - proj[:, 0].sum() / 4, (ny - 1) / 2 - proj[:, 1].sum() / 4]
    if flags & CALIB_ZERO_DISPARITY:
        cc[0][0] = cc[1][0] = (cc[0][0] + cc[1][0]) * 0.5
        cc[0][1] = cc[1][1] = (cc[0][1] + cc[1][1]) * 0.5
    elif idx == 0:
        cc[0][1] = cc[1][1] = (cc[0][1] + cc[1][1]) * 0.5
    else:
        cc[0][0] = cc[1][0] = (cc[0][0] + cc[1][0]) * 0.5
    P1 = np.zeros((3, 4))
    P1[0, 0] = P1[1, 1] = fc_new
    P1[0, 2], P1[1, 2], P1[2, 2] = cc[0][0], cc[0][1], 1
    P2 = P1.copy()
    P2[0, 2], P2[1, 2] = cc[1][0], cc[1][1]
    P2[idx, 3] = t[idx] * fc_new
    alpha = min(alpha, 1.0)
    inner1, outer1 = _rectangles(K[0], D[0], R1, P1, (nx, ny))
    inner2, outer2 = _rectangles(K[1], D[1], R2, P2, (nx, ny))
    cx1_0, cy1_0, cx2_0, cy2_0 = cc[0][0], cc[0][1], cc[1][0], cc[1][1]
    cx1, cy1 = new[0] * cx1_0 / nx, new[1] * cy1_0 / ny
    cx2, cy2 = new[0] * cx2_0 / nx, new[1] * cy2_0 / ny
    s = 1.0
    if alpha >= 0:
        def lims(cx, cy, cx0, cy0, r):
            return (cx / (cx0 - r[0]), cy / (cy0 - r[1]), (new[0] - 1 - cx) / (r[0] + r[2] - cx0),
                    (new[1] - 1 - cy) / (r[1] + r[3] - cy0))
        s0 = max(lims(cx1, cy1, cx1_0, cy1_0, inner1) + lims(cx2, cy2, cx2_0, cy2_0, inner2))
        s1 = min(lims(cx1, cy1, cx1_0, cy1_0, outer1) + lims(cx2, cy2, cx2_0, cy2_0, outer2))
        s = s0 * (1 - alpha) + s1 * alpha
    fc_new *= s
    for Pk, cx, cy in ((P1, cx1, cy1), (P2, cx2, cy2)):
        Pk[0, 0] = Pk[1, 1] = fc_new
        Pk[0, 2], Pk[1, 2] = cx, cy
    P2[idx, 3] = s * P2[idx, 3]

    def roi(inner, cx0, cy0, cx, cy):
        r = (int(np.ceil((inner[0] - cx0) * s + cx)), int(np.ceil((inner[1] - cy0) * s + cy)),
             int(np.floor(inner[2] * s)), int(np.floor(inner[3] * s)))
        return _rect_and(r, (0, 0, new[0], new[1]))
    roi1 = roi(inner1, cx1_0, cy1_0, cx1, cy1)
    roi2 = roi(inner2, cx2_0, cy2_0, cx2, cy2)
    Q = np.array([[1, 0, 0, -cx1], [0, 1, 0, -cy1], [0, 0, 0, fc_new],
                  [0, 0, -1.0 / t[idx], ((cx1 - cx2) if idx == 0 else (cy1 - cy2)) / t[idx]]])
    return R1, R2, P1, P2, Q, roi1, roi2


# ------------------------------------------------------------------------------------------------------------
# The reference's two loaders.  ROIs are stored by the reference but never read on its per-frame path; they equal
# cv2's except where an ROI edge lands exactly on a pixel boundary (the scale is chosen so that one does) and the
# last bit of a double decides the ceil (9 of 200 random calibrations: one row / column more or less).
# ------------------------------------------------------------------------------------------------------------
STEREO_CALIBRATION_FILE = os.path.join("output", "stereo_calibration_data.pkl")     # depth_map.py:22


def _read_pkl(path):
    with open(path, "rb") as f:
        data = pickle.load(f)
    missing = [k for k in CALIBRATION_KEYS if k not in data]
    if missing:
        raise KeyError(f"calibration file {path}: missing {missing}")
    return data


def _maps(K, dist, R, P, size, device_maps):
    from .cv_api import initUndistortRectifyMap, CV_16SC2
    return initUndistortRectifyMap(K, dist, R, P, size, CV_16SC2, as_tensor=bool(device_maps))


def load_stereo_calibration(path=None, device_maps=False):
    """depth_map.load_stereo_calibration (depth_map.py:591-668): the 7-key pickle -> stereoRectify(alpha=0) ->
    both CV_16SC2 map pairs -> the dict apply_stereo_rectification reads.  Returns None when the file is missing or
    unreadable, like the reference.  device_maps=True leaves the maps on the GPU as torch tensors (svb.remap takes
    them as they are, no per-frame upload and no map cache involved)."""
    path = STEREO_CALIBRATION_FILE if path is None else path
    if not os.path.exists(path):
        return None
    try:
        d = _read_pkl(path)
        img_size = tuple(int(v) for v in d["img_size"])
        T = np.asarray(d["T"], np.float64)
        R1, R2, P1, P2, Q, roi1, roi2 = stereoRectify(d["mtx_left"], d["dist_left"], d["mtx_right"], d["dist_right"],
                                                      img_size, d["R"], T, alpha=0)
        lm1, lm2 = _maps(d["mtx_left"], d["dist_left"], R1, P1, img_size, device_maps)
        rm1, rm2 = _maps(d["mtx_right"], d["dist_right"], R2, P2, img_size, device_maps)
        return {"left_map1": lm1, "left_map2": lm2, "right_map1": rm1, "right_map2": rm2, "roi1": roi1, "roi2": roi2,
                "Q": Q, "R": d["R"], "T": d["T"], "mtx_left": d["mtx_left"], "dist_left": d["dist_left"],
                "mtx_right": d["mtx_right"], "dist_right": d["dist_right"], "baseline": abs(T.reshape(3, -1)[0, 0]),
                "img_size": img_size}
    except Exception as e:                                      # the reference catches everything and returns None
        import warnings
        warnings.warn(f"load_stereo_calibration({path}): {e}")
        return None


def load_stereo_calibration_with_scaling(scale_factor=1.0, path=None, device_maps=False):
    """fused_depth_map.load_stereo_calibration_with_scaling (fused_depth_map.py:307-442): intrinsics scaled by
    scale_factor, image size int(W * s) x int(H * s), stereoRectify(alpha=0, CALIB_ZERO_DISPARITY), maps at the
    processing size.  The reference re-runs this when the user changes the scale ('+' / '-', :3339, :3355): with the
    maps made on the device that is a few hundred microseconds instead of two cv2 calls."""
    path = STEREO_CALIBRATION_FILE if path is None else path
    if not os.path.exists(path):
        return None
    try:
        d = _read_pkl(path)
        Kl = np.array(d["mtx_left"], np.float64)
        Kr = np.array(d["mtx_right"], np.float64)
        for K in (Kl, Kr):
            K[0, 0] *= scale_factor
            K[1, 1] *= scale_factor
            K[0, 2] *= scale_factor
            K[1, 2] *= scale_factor
        orig = tuple(int(v) for v in d["img_size"])
        size = (int(orig[0] * scale_factor), int(orig[1] * scale_factor))
        T = np.asarray(d["T"], np.float64)
        R1, R2, P1, P2, Q, roi1, roi2 = stereoRectify(Kl, d["dist_left"], Kr, d["dist_right"], size, d["R"], T,
                                                      alpha=0, flags=CALIB_ZERO_DISPARITY)
        lm1, lm2 = _maps(Kl, d["dist_left"], R1, P1, size, device_maps)
        rm1, rm2 = _maps(Kr, d["dist_right"], R2, P2, size, device_maps)
        return {"left_map1": lm1, "left_map2": lm2, "right_map1": rm1, "right_map2": rm2, "roi1": roi1, "roi2": roi2,
                "Q": Q, "R": d["R"], "T": d["T"], "mtx_left": Kl, "dist_left": d["dist_left"], "mtx_right": Kr,
                "dist_right": d["dist_right"], "baseline": abs(T.reshape(3, -1)[0, 0]), "img_size_orig": orig,
                "img_size_proc": size, "focal_length": Kl[0, 0], "scale_factor": scale_factor}
    except Exception as e:
        import warnings
        warnings.warn(f"load_stereo_calibration_with_scaling({path}): {e}")
        return None
