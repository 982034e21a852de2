"""CPU oracle for the data-parallel code around the mono estimators -- TEST INFRASTRUCTURE ONLY.

NumPy restatement of
    MiDaSDepthEstimator.estimate_depth after the network     fused_depth_map.py:1134-1158
    OpticalFlowDepthEstimator.estimate_depth after the flow  fused_depth_map.py:1386-1403
    _estimate_ego_motion's expected-flow grid                fused_depth_map.py:1491-1497
including the library calls they make: torch's upsample_bicubic2d (A = -0.75, align_corners=False),
cv2.GaussianBlur on CV_8U 5x5 sigma 0 (fixed-point binomial), cv2.normalize NORM_MINMAX on CV_8U and
cv2.perspectiveTransform on float32 points.  Pinned by tests/golden/golden_mono.npz, which
tests/golden/make_golden_mono.py produced with torch and cv2 themselves in the build container.
"""
import numpy as np

from .fusion import bilateral_filter

f32 = np.float32


def _reflect101(i, n):
    """cv2.borderInterpolate(i, n, BORDER_REFLECT_101), repeated reflection for images smaller than the kernel."""
    i = np.asarray(i).copy()
    if n == 1:
        return np.zeros_like(i)
    while ((i < 0) | (i >= n)).any():
        i = np.where(i < 0, -i, i)
        i = np.where(i >= n, 2 * (n - 1) - i, i)
    return i


def _cubic_weights(t, A=f32(-0.75)):
    t = t.astype(f32)
    one, two = f32(1), f32(2)
    x = t + one
    w0 = ((A * x - f32(5) * A) * x + f32(8) * A) * x - f32(4) * A
    w1 = ((A + two) * t - (A + f32(3))) * t * t + one
    x = one - t
    w2 = ((A + two) * x - (A + f32(3))) * x * x + one
    x = two - t
    w3 = ((A * x - f32(5) * A) * x + f32(8) * A) * x - f32(4) * A
    return [w.astype(f32) for w in (w0, w1, w2, w3)]


def resize_bicubic(src, Ho, Wo):
    """torch.nn.functional.interpolate(src[None, None], size=(Ho, Wo), mode="bicubic", align_corners=False)."""
    src = np.asarray(src, f32)
    Hi, Wi = src.shape
    sy, sx = f32(Hi) / f32(Ho), f32(Wi) / f32(Wo)
    ys = (np.arange(Ho, dtype=f32) + f32(0.5)) * sy - f32(0.5)
    xs = (np.arange(Wo, dtype=f32) + f32(0.5)) * sx - f32(0.5)
    iy, ix = np.floor(ys).astype(np.int64), np.floor(xs).astype(np.int64)
    wy, wx = _cubic_weights(ys - iy.astype(f32)), _cubic_weights(xs - ix.astype(f32))
    out = np.zeros((Ho, Wo), f32)
    for a in range(4):
        rows = src[np.clip(iy - 1 + a, 0, Hi - 1)]
        acc = np.zeros((Ho, Wo), f32)
        for c in range(4):
            acc += rows[:, np.clip(ix - 1 + c, 0, Wi - 1)] * wx[c][None, :]
        out += acc * wy[a][:, None]
    return out


def normalize_0_255(depth_raw):
    """fused_depth_map.py:1138-1145 (float32 throughout: NEP 50 keeps the Python scalars weak)."""
    d = np.asarray(depth_raw, f32)
    mn, mx = d.min(), d.max()
    if (mx - mn) < 1e-6:
        return np.zeros_like(d)
    return ((d - mn) / (mx - mn + f32(1e-8)) * f32(255.0)).astype(f32)


def gaussian_blur5_u8(gray):
    """cv2.GaussianBlur(gray u8, (5, 5), 0): (sum w_i w_j p + 128) >> 8 with w = {1,4,6,4,1}, REFLECT_101."""
    a = np.asarray(gray, np.uint8)
    H, W = a.shape
    k = np.array([1, 4, 6, 4, 1], np.int64)
    t = a.astype(np.int64)[:, _reflect101(np.arange(-2, W + 2), W)]
    row = sum(k[i] * t[:, i:i + W] for i in range(5))
    t = row[_reflect101(np.arange(-2, H + 2), H), :]
    s = sum(k[i] * t[i:i + H, :] for i in range(5))
    return ((s + 128) >> 8).astype(np.uint8)


def normalize_minmax_u8(c, lo=0.0, hi=1.0):
    """cv2.normalize(c u8, None, lo, hi, cv2.NORM_MINMAX): scale / shift in double, convertTo in float32."""
    c = np.asarray(c, np.uint8)
    mn, mx = int(c.min()), int(c.max())
    scale = (hi - lo) * (1.0 / (mx - mn) if mx - mn > np.finfo(np.float64).eps else 0.0)
    shift = lo - mn * scale
    v = c.astype(f32) * f32(scale) + f32(shift)
    return np.clip(np.rint(v), 0, 255).astype(np.uint8)


def contrast_confidence(gray):
    """fused_depth_map.py:1150-1156."""
    g = np.asarray(gray, np.uint8)
    b = gaussian_blur5_u8(g)
    c = np.abs(g.astype(np.int16) - b.astype(np.int16)).astype(np.uint8)
    return normalize_minmax_u8(c)


def midas_postprocess(prediction, gray):
    """-> (depth_normalized f32, confidence u8) for one frame."""
    H, W = np.asarray(gray).shape
    return normalize_0_255(resize_bicubic(prediction, H, W)), contrast_confidence(gray)


def expected_flow(Hm, h, w):
    """cv2.perspectiveTransform(grid f32, H) - grid (fused_depth_map.py:1491-1497)."""
    m = np.asarray(Hm, np.float64).ravel()
    xs, ys = np.meshgrid(np.arange(w, dtype=f32), np.arange(h, dtype=f32))
    x, y = xs.astype(np.float64), ys.astype(np.float64)
    wv = x * m[6] + y * m[7] + m[8]
    ok = np.abs(wv) > np.finfo(np.float64).eps
    inv = np.where(ok, 1.0 / np.where(ok, wv, 1.0), 0.0)
    wx = ((x * m[0] + y * m[1] + m[2]) * inv).astype(f32)
    wy = ((x * m[3] + y * m[4] + m[5]) * inv).astype(f32)
    return np.stack([wx - xs, wy - ys], -1)


def flow_residual_depth(flow, Hm, stable_depth=None, camera_moving=True):
    """fused_depth_map.py:1386-1403 -> (depth_smoothed, stable_depth, current_depth)."""
    flow = np.asarray(flow, f32)
    h, w = flow.shape[:2]
    ef = expected_flow(Hm, h, w)
    rx, ry = flow[..., 0] - ef[..., 0], flow[..., 1] - ef[..., 1]
    mag = np.hypot(rx, ry)
    cur = (1.0 / (mag + 0.5)).astype(f32)
    if stable_depth is None:
        stable = cur.copy()
    else:
        alpha = 0.9 if camera_moving else 0.99
        stable = (alpha * np.asarray(stable_depth, f32) + (1 - alpha) * cur).astype(f32)
    return bilateral_filter(cur, 9, 75.0, 75.0), stable, cur
