"""CPU oracle for cv2.initUndistortRectifyMap(K, dist, R, P, size, CV_16SC2) -- TEST INFRASTRUCTURE ONLY.

NumPy float64 restatement of OpenCV 4.13's imgproc/src/undistort.dispatch.cpp (the call the reference makes
at depth_map.py:633-638 and fused_depth_map.py:404-409): back-project every destination pixel through
inv(P[:, :3] @ R), apply the rational + tangential + thin-prism distortion model, project with K, and store
the position in the fixed-point form cv2.remap consumes (INTER_BITS = 5: map1 = integer part, map2 = 32*fy+fx).
Pinned against live cv2 by tests/test_oracle_golden.py and against tests/golden/golden.npz.
"""
import numpy as np

INTER_BITS = 5
INTER_TAB_SIZE = 1 << INTER_BITS


def init_undistort_rectify_map(K, dist, R, P, size):
    W, H = size
    K = np.asarray(K, np.float64)
    d = np.zeros(14, np.float64)
    dist = np.asarray(dist, np.float64).ravel() if dist is not None else np.zeros(0)
    d[:dist.size] = dist
    k1, k2, p1, p2, k3, k4, k5, k6, s1, s2, s3, s4, tx, ty = d
    if tx != 0 or ty != 0:
        raise NotImplementedError("tilted sensor model")
    R = np.eye(3) if R is None else np.asarray(R, np.float64)
    A = np.asarray(P, np.float64)[:3, :3]
    ir = np.linalg.inv(A @ R)
    fx, fy, u0, v0 = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    j = np.arange(W, dtype=np.float64)[None, :]
    i = np.arange(H, dtype=np.float64)[:, None]
    _x = j * ir[0, 0] + (i * ir[0, 1] + ir[0, 2])
    _y = j * ir[1, 0] + (i * ir[1, 1] + ir[1, 2])
    _w = j * ir[2, 0] + (i * ir[2, 1] + ir[2, 2])
    w = 1.0 / _w
    x, y = _x * w, _y * w
    x2, y2 = x * x, y * y
    r2 = x2 + y2
    _2xy = 2 * x * y
    kr = (1 + ((k3 * r2 + k2) * r2 + k1) * r2) / (1 + ((k6 * r2 + k5) * r2 + k4) * r2)
    xd = x * kr + p1 * _2xy + p2 * (r2 + 2 * x2) + s1 * r2 + s2 * r2 * r2
    yd = y * kr + p1 * (r2 + 2 * y2) + p2 * _2xy + s3 * r2 + s4 * r2 * r2
    u = fx * xd + u0
    v = fy * yd + v0
    iu = np.clip(np.rint(u * INTER_TAB_SIZE), -2.0 ** 31, 2.0 ** 31 - 1).astype(np.int64)   # saturate_cast<int>(cvRound)
    iv = np.clip(np.rint(v * INTER_TAB_SIZE), -2.0 ** 31, 2.0 ** 31 - 1).astype(np.int64)
    map1 = np.stack([iu >> INTER_BITS, iv >> INTER_BITS], -1).astype(np.int16)              # (short) truncation wraps
    map2 = ((iv & (INTER_TAB_SIZE - 1)) * INTER_TAB_SIZE + (iu & (INTER_TAB_SIZE - 1))).astype(np.uint16)
    return map1, map2
