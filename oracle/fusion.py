"""CPU oracle for the fusion chain of fused_depth_map.py -- TEST INFRASTRUCTURE ONLY.

NumPy restatement (float32 where the reference's NumPy expressions stay in float32, NEP 50) of
    create_depth_map_stereo_scaled tail   fused_depth_map.py:1010-1029
    calibrate_midas_to_stereo             fused_depth_map.py:1169-1257
    normalize_to_stereo_range             fused_depth_map.py:1503-1554
    fuse_depth_maps                       fused_depth_map.py:1560-1698
    detect_camera_occlusion               fused_depth_map.py:131-301
including the two OpenCV filters they call (GaussianBlur 15x15 sigma 0 and bilateralFilter
9/75/75 on float32; opencv 4.13 behaviour, pinned numerically against cv2 by
tests/golden/make_golden_fusion.py).  Parity for this file is pinned by the golden vectors that
script produced by importing the reference's own functions from /root/reference.
"""
import numpy as np

f32 = np.float32

DEFAULT_FUSION_PARAMS = dict(stereo_weight=0.8, midas_fill_weight=0.9, flow_fill_weight=0.5,
                             stereo_conf_threshold=0.5, flow_hole_threshold=15)   # fused_depth_map.py:81-116


def normalize_disparity(disparity, min_disp, num_disp):
    """fused_depth_map.py:1010-1029 -> (normalized f32 0..254, confidence f32)."""
    d = np.asarray(disparity, f32)
    clipped = np.clip(d, min_disp, min_disp + num_disp - 1)
    norm = ((clipped - f32(min_disp)) / f32(num_disp) * f32(255.0)).astype(np.uint8).astype(f32)
    conf = ((d > min_disp + 1) & (d < min_disp + num_disp - 1)).astype(f32)
    return norm, conf


def percentile(values, q, scalar_q=True):
    """np.percentile(values, q) with linear interpolation on a flat float32 array (NumPy 2.x):
    a Python-scalar q is divided by float32(100) and the virtual index, the weight and the lerp
    stay in float32; a list q ([10, 90]) becomes float64 and so does everything after it."""
    a = np.sort(np.asarray(values, f32).ravel())
    n = a.size
    T = f32 if scalar_q else np.float64
    qq = T(q) / T(100)
    vi = T(T(n - 1) * qq)                                            # method 'linear': (n - 1) * q
    if vi >= n - 1:
        return T(a[-1])
    lo = int(np.floor(vi)) if vi >= 0 else 0
    hi = lo + 1
    g = T(vi - T(lo)) if vi >= 0 else T(0)
    x, y = a[lo], a[hi]
    d = y - x                                                        # float32 subtract
    # _lerp: a + d*t, replaced by b - d*(1-t) where t >= 0.5
    r = T(x) + T(d) * g if g < 0.5 else T(y) - T(d) * (T(1) - g)
    return T(r)


def calibrate_midas_to_stereo(midas, stereo_disparity, stereo_conf):
    """fused_depth_map.py:1169-1257, including the cv2.resize(INTER_LINEAR) of a mono map of another shape
    (:1214-1217; oracle.resize restates it)."""
    m = np.asarray(midas, f32)
    s = np.asarray(stereo_disparity, f32)
    if m.shape != s.shape:
        from . import resize
        m = resize(np.ascontiguousarray(m), (s.shape[1], s.shape[0]))
    rel = np.asarray(stereo_conf) > 0.7
    if int(rel.sum()) < 100:
        mmin, mmax = percentile(m, 5), percentile(m, 95)
        smin, smax = percentile(s, 5), percentile(s, 95)
        if (mmax - mmin) < 1e-6:
            return np.full_like(m, (smin + smax) / f32(2.0))
        norm = (m - mmin) / (mmax - mmin + f32(1e-8))
        return (smin + norm * (smax - smin)).astype(f32)
    sv, mv = s[rel], m[rel]
    smin, smax = percentile(sv, 10, False), percentile(sv, 90, False)     # np.percentile(v, [10, 90]): float64
    mmin, mmax = percentile(mv, 10, False), percentile(mv, 90, False)
    scale = 1.0 if (mmax - mmin) < 1e-6 else (smax - smin) / (mmax - mmin + 1e-8)
    offset = smin - mmin * scale
    return (m.astype(np.float64) * scale + offset).astype(f32)


def normalize_to_stereo_range(depth_map, stereo_disparity):
    """fused_depth_map.py:1503-1554."""
    d = np.asarray(depth_map, f32)
    s = np.asarray(stereo_disparity, f32)
    valid = s > 0
    if valid.any():
        smin, smax = percentile(s[valid], 5), percentile(s[valid], 95)
    else:
        smin, smax = f32(0), f32(255)
    dmin, dmax = percentile(d, 5), percentile(d, 95)
    if (dmax - dmin) < 1e-6:
        return np.full_like(d, (smin + smax) / f32(2.0))
    norm = (d - dmin) / (dmax - dmin + f32(1e-8))
    return (smin + norm * (smax - smin)).astype(f32)


def _reflect101(i, n):
    i = np.abs(i)
    return np.where(i >= n, 2 * (n - 1) - i, i)


def gaussian_kernel(ksize):
    """cv2.getGaussianKernel(ksize, sigma<=0) as float32."""
    sigma = 0.3 * ((ksize - 1) * 0.5 - 1) + 0.8
    x = np.arange(ksize) - (ksize - 1) / 2.0
    k = np.exp(-x * x / (2 * sigma * sigma))
    return (k / k.sum()).astype(f32)


def gaussian_blur(a, ksize=15):
    """cv2.GaussianBlur(a, (ksize, ksize), 0) on float32, BORDER_REFLECT_101 (rows, then columns)."""
    a = np.asarray(a, f32)
    H, W = a.shape
    k = gaussian_kernel(ksize)
    r = ksize // 2
    xi = _reflect101(np.arange(-r, W + r), W)
    t = a[:, xi]
    row = np.zeros((H, W), f32)
    for i in range(ksize):
        row += k[i] * t[:, i:i + W]
    yi = _reflect101(np.arange(-r, H + r), H)
    t = row[yi, :]
    out = np.zeros((H, W), f32)
    for i in range(ksize):
        out += k[i] * t[i:i + H, :]
    return out


def bilateral_filter(src, d=9, sigma_color=75.0, sigma_space=75.0):
    """cv2.bilateralFilter on single-channel float32 (opencv 4.13: 4096-bin exp table with linear
    interpolation, circular window, BORDER_REFLECT_101)."""
    src = np.asarray(src, f32)
    H, W = src.shape
    r = d // 2
    gc = -0.5 / (sigma_color * sigma_color)
    gs = -0.5 / (sigma_space * sigma_space)
    mn, mx = float(src.min()), float(src.max())
    ln = f32(mx - mn)
    if ln <= np.finfo(f32).eps:
        return src.copy()
    nb = 1 << 12
    scale = f32(nb) / ln
    lut = np.zeros(nb + 2, f32)
    last = 1.0
    for i in range(nb + 2):
        if last > 0:
            v = i / float(scale)
            lut[i] = f32(np.exp(v * v * gc))
            last = lut[i]
    yi = _reflect101(np.arange(-r, H + r), H)
    xi = _reflect101(np.arange(-r, W + r), W)
    tmp = src[yi][:, xi]
    s = src.copy()
    ws = np.ones((H, W), f32)
    for i in range(-r, r + 1):
        for j in range(-r, r + 1):
            rr = np.sqrt(float(i * i + j * j))
            if rr > r or (i == 0 and j == 0):
                continue
            sw = f32(np.exp(rr * rr * gs))
            val = tmp[r + i:r + i + H, r + j:r + j + W]
            al = np.abs(val - src) * scale
            idx = np.floor(al).astype(np.int32)
            al = (al - idx.astype(f32)).astype(f32)
            w = sw * (lut[idx] + al * (lut[idx + 1] - lut[idx]))
            s += val * w
            ws += w
    return (s / ws).astype(f32)


def fuse_depth_maps(stereo_depth, stereo_conf, midas_cal, midas_conf, flow_norm, camera_moving,
                    use_stereo=True, use_midas=True, use_flow=True, fusion_params=None,
                    use_midas_fill=True, use_flow_fill=True, blend_radius=15, bilateral=(9, 75, 75)):
    """fused_depth_map.py:1560-1698 -> (fused_uint8, fused_f32_before_cast) or None."""
    P = dict(DEFAULT_FUSION_PARAMS)
    if fusion_params:
        P.update(fusion_params)
    vs = use_stereo and stereo_depth is not None and np.size(stereo_depth) > 0
    vm = use_midas and midas_cal is not None and np.size(midas_cal) > 0
    vf = use_flow and flow_norm is not None and np.size(flow_norm) > 0 and camera_moving
    if not (vs or vm or vf):
        return None
    if vs:
        fused = np.asarray(stereo_depth, f32).copy()
        conf = np.asarray(stereo_conf, f32).copy() if stereo_conf is not None else np.ones_like(fused)
        fused *= f32(P['stereo_weight'])
        if vm and use_midas_fill:
            low = conf < P['stereo_conf_threshold']
            if low.any():
                fw = (f32(1.0) - conf) * f32(P['midas_fill_weight'])
                fw = np.clip(gaussian_blur(fw, blend_radius), 0.0, 1.0).astype(f32)
                mask = low & (fw > 0.1)
                if mask.any():
                    fused[mask] = fused[mask] * (1 - fw[mask]) + np.asarray(midas_cal, f32)[mask] * fw[mask]
                    conf[mask] = np.maximum(conf[mask], 0.8)
        if vf and use_flow_fill:
            hole = (fused < P['flow_hole_threshold']) | (fused == 0)
            if hole.any():
                w = P['flow_fill_weight']
                fused[hole] = fused[hole] * f32(1 - w) + np.asarray(flow_norm, f32)[hole] * f32(w)
    elif vm:
        fused = np.asarray(midas_cal, f32).copy()
        if vf and use_flow_fill:
            hole = (fused < P['flow_hole_threshold']) | (fused == 0)
            if hole.any():
                w = P['flow_fill_weight']
                fused[hole] = fused[hole] * f32(1 - w) + np.asarray(flow_norm, f32)[hole] * f32(w)
    else:
        fused = np.asarray(flow_norm, f32).copy()
    if fused.max() > 10.0:
        fused = bilateral_filter(fused, *bilateral)
    fused = np.clip(fused, 0, 255)
    return fused.astype(np.uint8), fused


def occlusion_stats(gray, block=48):
    """The five statistics of fused_depth_map.py:170-240 for one camera:
    (mean block std, low-variance block ratio, global std, histogram entropy, mean)."""
    g = np.asarray(gray)
    h, w = g.shape
    bh, bw = max(1, h // block), max(1, w // block)
    stds = []
    for i in range(bh):
        for j in range(bw):
            blk = g[i * block:min((i + 1) * block, h), j * block:min((j + 1) * block, w)]
            if blk.size:
                stds.append(np.std(blk))
    avg = float(np.mean(stds)) if stds else 0.0
    low = float(np.sum(np.array(stds) < 12) / max(len(stds), 1))
    hist = np.bincount(g.ravel(), minlength=256).astype(np.float32) + 1e-10
    hist = hist / hist.sum()
    ent = float(-np.sum(hist * np.log2(hist + 1e-10)))
    return avg, low, float(np.std(g)), ent, float(np.mean(g))


def occlusion_decision(left_stats, right_stats, threshold=0.45):
    """Penalty scores and decision of fused_depth_map.py:242-301."""
    def score(a, b):
        s = 0.0
        if a[0] < 28.0 * 0.8:
            s += 0.35
        if a[1] > 0.55:
            s += 0.35
        if a[2] < b[2] / 2.2 and b[2] > 15:
            s += 0.25
        if a[3] < b[3] / 1.6 and b[3] > 5.0:
            s += 0.25
        if abs(a[4] - b[4]) > 45.0 and a[4] < 80:
            s += 0.2
        return s
    ls, rs = score(left_stats, right_stats), score(right_stats, left_stats)
    if ls > threshold and rs < threshold * 0.6:
        res = 'left'
    elif rs > threshold and ls < threshold * 0.6:
        res = 'right'
    elif ls > threshold and rs > threshold:
        res = 'both'
    else:
        res = 'none'
    return res, ls, rs


def detect_camera_occlusion(gray_left, gray_right, threshold=0.45):
    return occlusion_decision(occlusion_stats(gray_left), occlusion_stats(gray_right), threshold)
