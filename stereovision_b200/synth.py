"""Deterministic synthetic stereo pairs: no webcams on the box.

NumPy only.  The scene is a textured background plane plus a few fronto-parallel
rectangles painted far-to-near, so occlusions are physically consistent; the right
view is the left texture shifted by the per-layer disparity, plus integer noise.
"""
import numpy as np


def _gauss_blur(a, sigma):
    r = max(1, int(3 * sigma + 0.5))
    k = np.exp(-0.5 * (np.arange(-r, r + 1) / sigma) ** 2)
    k /= k.sum()
    a = np.pad(a, ((r, r), (r, r)), mode="reflect")
    a = np.apply_along_axis(lambda v: np.convolve(v, k, mode="valid"), 0, a)
    a = np.apply_along_axis(lambda v: np.convolve(v, k, mode="valid"), 1, a)
    return a


def make_pair(H, W, D, seed=0, noise=6):
    """Returns (left u8 [H,W], right u8 [H,W], true_disparity int32 [H,W])."""
    rng = np.random.default_rng(seed)
    pad = 2 * D
    tex = _gauss_blur(rng.uniform(0, 255, (H, W + pad)), 1.2)
    tex = ((tex - tex.min()) / (tex.max() - tex.min()) * 255).astype(np.uint8)
    dB = int(rng.integers(2, max(3, D // 8) + 1))
    layers = [(0, H, 0, W, dB, tex)]
    for _ in range(int(rng.integers(3, 7))):
        d = int(rng.integers(max(dB + 1, D // 8), max(dB + 2, int(0.8 * D))))
        h, w = int(rng.integers(H // 6, H // 2)), int(rng.integers(W // 8, W // 2))
        y0, x0 = int(rng.integers(0, H - h)), int(rng.integers(0, W - w))
        t = _gauss_blur(rng.uniform(0, 255, (H, W + pad)), 1.2)
        t = ((t - t.min()) / (t.max() - t.min()) * 255).astype(np.uint8)
        layers.append((y0, y0 + h, x0, x0 + w, d, t))
    layers.sort(key=lambda l: l[4])
    left = np.zeros((H, W), np.uint8)
    right = np.zeros((H, W), np.uint8)
    disp = np.zeros((H, W), np.int32)
    for (y0, y1, x0, x1, d, t) in layers:
        left[y0:y1, x0:x1] = t[y0:y1, pad // 2 + x0: pad // 2 + x1]
        disp[y0:y1, x0:x1] = d
        xr0, xr1 = max(0, x0 - d), max(0, x1 - d)
        if xr1 > xr0:
            right[y0:y1, xr0:xr1] = t[y0:y1, pad // 2 + xr0 + d: pad // 2 + xr1 + d]
    if W > dB:
        right[:, W - dB:] = tex[:, pad // 2 + W: pad // 2 + W + dB]
    if noise:
        n = rng.integers(-noise, noise + 1, (H, W))
        right = np.clip(right.astype(np.int32) + n, 0, 255).astype(np.uint8)
    return left, right, disp


def make_batch(B, H, W, D, seed0=0):
    L = np.empty((B, H, W), np.uint8)
    R = np.empty((B, H, W), np.uint8)
    for i in range(B):
        L[i], R[i], _ = make_pair(H, W, D, seed0 + i)
    return L, R


def to_bgr(gray):
    """gray replicated x3 with per-channel gain 0.9/1.0/1.1."""
    g = gray.astype(np.float32)
    return np.clip(np.stack([g * 0.9, g, g * 1.1], -1), 0, 255).astype(np.uint8)


def make_aux_depths(disp_true, seed=0):
    """Stand-ins for the two learned/flow inputs of the fusion chain (fused_depth_map.py:1560):
    a MiDaS-like relative inverse depth (smooth, arbitrary scale and offset, float32) and a
    flow-derived depth already scaled to 0..255 (float32, noisy, with zero patches)."""
    rng = np.random.default_rng(1000 + seed)
    d = disp_true.astype(np.float64)
    H, W = d.shape
    midas = _gauss_blur(d, 2.0) * rng.uniform(20.0, 60.0) + rng.uniform(100.0, 900.0)
    midas += _gauss_blur(rng.normal(0, 1, (H, W)), 3.0) * 40.0
    flow = _gauss_blur(d, 4.0)
    flow = (flow - flow.min()) / max(flow.max() - flow.min(), 1e-9) * 255.0
    flow += rng.normal(0, 4.0, (H, W))
    for _ in range(3):
        y0, x0 = int(rng.integers(0, H - H // 6)), int(rng.integers(0, W - W // 6))
        flow[y0:y0 + H // 6, x0:x0 + W // 6] = 0.0
    return midas.astype(np.float32), np.clip(flow, 0, 255).astype(np.float32)


def synthetic_calibration(W, H):
    """The synthetic stereo calibration of SURVEY 8(d) as the 7 keys the reference's loaders read
    (depth_map.py:613-619): fx = fy = 0.78 W, slightly off-centre principal point, mild radial / tangential
    distortion, a small relative rotation, T = 6 cm along -x (the reference's unit is centimetres)."""
    from .calib import Rodrigues
    fx = 0.78 * W
    K = np.array([[fx, 0, W / 2 + 3], [0, fx, H / 2 - 2], [0, 0, 1]], np.float64)
    return dict(mtx_left=K.copy(), dist_left=np.array([[-0.10, 0.02, 1e-3, -1e-3, 0]], np.float64),
                mtx_right=K.copy(), dist_right=np.array([[-0.09, 0.015, -1e-3, 1e-3, 0]], np.float64),
                R=Rodrigues(np.array([0.01, -0.02, 0.005])), T=np.array([[-6.0], [0.05], [0.02]]), img_size=(W, H))


def write_calibration_pkl(path, W, H):
    """Write the synthetic calibration where the reference looks for it (output/stereo_calibration_data.pkl,
    depth_map.py:22) -- BASELINE.json configs[0] 'synthetic calib .pkl'.  Returns the dict written."""
    import os
    import pickle
    d = synthetic_calibration(W, H)
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "wb") as f:
        pickle.dump(d, f)
    return d
