"""The fusion chain of fused_depth_map.py behind the reference's own function names, running on
libsvb.so (csrc/fuse.cu).  NumPy in -> NumPy out, CUDA tensors in -> CUDA tensors out (no stream
synchronisation); a leading batch dimension is accepted everywhere.

    normalize_disparity          fused_depth_map.py:1010-1029  (tail of create_depth_map_stereo_scaled)
    calibrate_midas_to_stereo    fused_depth_map.py:1169-1257
    normalize_to_stereo_range    fused_depth_map.py:1503-1554
    fuse_depth_maps              fused_depth_map.py:1560-1698  (the JET colormap / putText overlay is display glue)
    detect_camera_occlusion      fused_depth_map.py:131-301
"""
import ctypes

import numpy as np

from . import _lib
from .cv_api import _ensure_device, _restores_device, _is_tensor, _ret, _stream_ptr, _to_dev, _torch, cvtColor, resize

# fused_depth_map.py:80-116
FUSION_WEIGHTS = {"stereo_base": 0.8, "midas_max_fill": 0.9, "flow_max_fill": 0.5}
FUSION_THRESHOLDS = {"stereo_low_conf": 0.5, "midas_fill_start": 0.3, "flow_hole_threshold": 15, "flow_min_inliers": 15}
FUSION_SMOOTHING = {"midas_blend_radius": 15, "fusion_bilateral_d": 9, "fusion_bilateral_sigma": 75}
FUSION_METHODS = {"use_stereo": True, "use_midas_fill": True, "use_flow_fill": True}
fusion_params = {
    "stereo_weight": FUSION_WEIGHTS["stereo_base"],
    "midas_fill_weight": FUSION_WEIGHTS["midas_max_fill"],
    "flow_fill_weight": FUSION_WEIGHTS["flow_max_fill"],
    "stereo_conf_threshold": FUSION_THRESHOLDS["stereo_low_conf"],
    "flow_hole_threshold": FUSION_THRESHOLDS["flow_hole_threshold"],
}


def _f32_maps(*arrays):
    """-> (device tensors (None kept), was_tensor, B, H, W, batched); all maps must share a shape."""
    torch = _torch()
    out, was, shape = [], False, None
    for a in arrays:
        if a is None:
            out.append(None)
            continue
        t, w = _to_dev(a, np.float32) if not _is_tensor(a) else _to_dev(a)
        if t.dtype != torch.float32:
            t = t.float()
        was = was or w
        if shape is None:
            shape = tuple(t.shape)
        elif tuple(t.shape) != shape:
            raise ValueError(f"shape mismatch: {tuple(t.shape)} vs {shape}")
        out.append(t)
    if shape is None or len(shape) not in (2, 3):
        raise ValueError("expected [H,W] or [B,H,W] maps")
    B = 1 if len(shape) == 2 else int(shape[0])
    return out, was, B, int(shape[-2]), int(shape[-1])


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _scratch(nbytes, device):
    return _torch().empty(int(nbytes), dtype=_torch().uint8, device=device)


@_restores_device
def normalize_disparity(disparity, min_disp, num_disp):
    """(disparity_normalized f32 in 0..254, confidence f32 0/1) -- fused_depth_map.py:1010-1029."""
    torch = _torch()
    (d,), was, B, H, W = _f32_maps(disparity)
    _ensure_device(d.device.index)
    norm, conf = torch.empty_like(d), torch.empty_like(d)
    _lib.check(_lib.lib().svb_normalize_disparity_dev(d.data_ptr(), B, H, W, int(min_disp), int(num_disp),
                                                      norm.data_ptr(), conf.data_ptr(), _stream_ptr()))
    return _ret(norm, was), _ret(conf, was)


@_restores_device
def calibrate_midas_to_stereo(midas_depth, stereo_disparity, stereo_confidence):
    """fused_depth_map.py:1169-1257, including its first step: a mono map of another shape is resized to the
    stereo shape with cv2.resize(INTER_LINEAR) (:1214-1217) -- on the device (svb.resize, float32)."""
    if midas_depth is None or stereo_disparity is None:
        return None
    torch = _torch()
    if tuple(midas_depth.shape) != tuple(stereo_disparity.shape):
        if len(midas_depth.shape) != len(stereo_disparity.shape):
            raise ValueError(f"shape mismatch: {tuple(midas_depth.shape)} vs {tuple(stereo_disparity.shape)}")
        if not _is_tensor(midas_depth):
            midas_depth = np.asarray(midas_depth, np.float32)
        elif midas_depth.dtype != torch.float32:
            midas_depth = midas_depth.float()
        midas_depth = resize(midas_depth, (int(stereo_disparity.shape[-1]), int(stereo_disparity.shape[-2])))
    (m, s, c), was, B, H, W = _f32_maps(midas_depth, stereo_disparity, stereo_confidence)
    _ensure_device(m.device.index)
    L = _lib.lib()
    out = torch.empty_like(m)
    nb = L.svb_range_scratch_bytes(B)
    scr = _scratch(nb, m.device)
    _lib.check(L.svb_calibrate_midas_dev(m.data_ptr(), s.data_ptr(), c.data_ptr(), B, H, W, out.data_ptr(),
                                         scr.data_ptr(), nb, _stream_ptr()))
    return _ret(out, was)


@_restores_device
def normalize_to_stereo_range(depth_map, stereo_disparity):
    """fused_depth_map.py:1503-1554."""
    if depth_map is None or stereo_disparity is None:
        return None
    torch = _torch()
    (d, s), was, B, H, W = _f32_maps(depth_map, stereo_disparity)
    _ensure_device(d.device.index)
    L = _lib.lib()
    out = torch.empty_like(d)
    nb = L.svb_range_scratch_bytes(B)
    scr = _scratch(nb, d.device)
    _lib.check(L.svb_normalize_to_stereo_range_dev(d.data_ptr(), s.data_ptr(), B, H, W, out.data_ptr(),
                                                   scr.data_ptr(), nb, _stream_ptr()))
    return _ret(out, was)


@_restores_device
def percentiles(values, mask=None, mask_thr=0.0, mask_mode=0):
    """np.percentile(values[mask], 5 / 95) per image (mask_mode as in include/svb.h).
    Returns (float64 [B,2], counts uint32 [B])."""
    torch = _torch()
    (v, m), was, B, H, W = _f32_maps(values, mask)
    _ensure_device(v.device.index)
    L = _lib.lib()
    out = torch.empty((B, 2), dtype=torch.float64, device=v.device)
    cnt = torch.empty(B, dtype=torch.int32, device=v.device)
    nb = L.svb_range_scratch_bytes(B)
    scr = _scratch(nb, v.device)
    _lib.check(L.svb_percentiles_f32_dev(v.data_ptr(), _ptr(m), float(mask_thr), int(mask_mode), B, H * W,
                                         out.data_ptr(), cnt.data_ptr(), scr.data_ptr(), nb, _stream_ptr()))
    return _ret(out, was), _ret(cnt, was)


@_restores_device
def fuse_depth_maps(stereo_depth, stereo_conf, midas_depth_calibrated, midas_conf, flow_depth_normalized,
                    camera_moving, use_stereo=True, use_midas=True, use_flow=True, fusion_params=None,
                    return_float=False):
    """fused_depth_map.py:1560-1698.  Returns the fused uint8 map (the reference's second return value,
    a JET colormap with text drawn on it, is display glue); None when no method is valid.
    `fusion_params` overrides this module's `fusion_params` (the reference's module global);
    `return_float=True` also returns the float32 map before the uint8 cast."""
    torch = _torch()
    P = dict(globals()["fusion_params"])
    if fusion_params:
        P.update(fusion_params)

    def size(a):
        return 0 if a is None else (a.numel() if _is_tensor(a) else np.size(a))
    valid_stereo = bool(use_stereo) and size(stereo_depth) > 0
    valid_midas = bool(use_midas) and size(midas_depth_calibrated) > 0
    valid_flow = bool(use_flow) and size(flow_depth_normalized) > 0 and bool(camera_moving)
    if not (valid_stereo or valid_midas or valid_flow):
        return None
    (s, c, m, f), was, B, H, W = _f32_maps(stereo_depth if valid_stereo else None,
                                           stereo_conf if valid_stereo else None,
                                           midas_depth_calibrated if valid_midas else None,
                                           flow_depth_normalized if valid_flow else None)
    first = next(t for t in (s, m, f) if t is not None)
    _ensure_device(first.device.index)
    L = _lib.lib()
    prm = _lib.FuseParams(float(P["stereo_weight"]), float(P["midas_fill_weight"]), float(P["flow_fill_weight"]),
                          float(P["stereo_conf_threshold"]), float(P["flow_hole_threshold"]),
                          float(FUSION_SMOOTHING["fusion_bilateral_sigma"]), float(FUSION_SMOOTHING["fusion_bilateral_sigma"]),
                          int(FUSION_SMOOTHING["midas_blend_radius"]), int(FUSION_SMOOTHING["fusion_bilateral_d"]),
                          int(bool(FUSION_METHODS["use_midas_fill"])), int(bool(FUSION_METHODS["use_flow_fill"])))
    out8 = torch.empty(first.shape, dtype=torch.uint8, device=first.device)
    outf = torch.empty_like(first) if return_float else None
    nb = L.svb_fuse_scratch_bytes(B, H, W)
    scr = _scratch(nb, first.device)
    _lib.check(L.svb_fuse_depth_maps_dev(_ptr(s), _ptr(c), _ptr(m), _ptr(f), B, H, W, ctypes.byref(prm),
                                         out8.data_ptr(), _ptr(outf), scr.data_ptr(), nb, _stream_ptr()))
    if return_float:
        return _ret(out8, was), _ret(outf, was)
    return _ret(out8, was)


@_restores_device
def occlusion_stats(gray, block_size=48):
    """Per image: (mean block std, low-variance block ratio, global std, entropy, mean) --
    fused_depth_map.py:170-240.  gray: u8 [H,W] or [B,H,W]; returns float64 [5] or [B,5]."""
    torch = _torch()
    g, was = _to_dev(gray)
    if g.dtype != torch.uint8 or g.ndim not in (2, 3):
        raise ValueError("expected uint8 [H,W] or [B,H,W]")
    B = 1 if g.ndim == 2 else int(g.shape[0])
    H, W = int(g.shape[-2]), int(g.shape[-1])
    _ensure_device(g.device.index)
    L = _lib.lib()
    stats = torch.empty((B, 5), dtype=torch.float64, device=g.device)
    nb = L.svb_occlusion_scratch_bytes(B, H, W, int(block_size))
    scr = _scratch(nb, g.device)
    _lib.check(L.svb_occlusion_stats_u8_dev(g.data_ptr(), B, H, W, int(block_size), stats.data_ptr(),
                                            scr.data_ptr(), nb, _stream_ptr()))
    if g.ndim == 2:
        stats = stats[0]
    return _ret(stats, was)


def detect_camera_occlusion(left_img, right_img, occlusion_threshold=0.45):
    """fused_depth_map.py:131-301 -> (result, left_score, right_score), result in
    'none' / 'left' / 'right' / 'both'.  The image statistics run on the device (both cameras in one
    batch); the five threshold tests on ten scalars are host logic like the reference's."""
    torch = _torch()

    def gray(img):
        t, _ = _to_dev(img)
        return cvtColor(t) if t.ndim == 3 and t.shape[-1] == 3 else t
    gl, gr = gray(left_img), gray(right_img)
    if gl.shape == gr.shape:
        st = occlusion_stats(torch.stack([gl, gr])).cpu().numpy()
        l, r = st[0], st[1]
    else:
        l, r = occlusion_stats(gl).cpu().numpy(), occlusion_stats(gr).cpu().numpy()
    STD_THRESHOLD, LOW_VAR_THRESHOLD, CONTRAST_RATIO, ENTROPY_RATIO, BRIGHTNESS_DIFF = 28.0, 0.55, 2.2, 1.6, 45.0

    def score(a, b):
        s = 0.0
        if a[0] < STD_THRESHOLD * 0.8:
            s += 0.35
        if a[1] > LOW_VAR_THRESHOLD:
            s += 0.35
        if a[2] < b[2] / CONTRAST_RATIO and b[2] > 15:
            s += 0.25
        if a[3] < b[3] / ENTROPY_RATIO and b[3] > 5.0:
            s += 0.25
        if abs(a[4] - b[4]) > BRIGHTNESS_DIFF and a[4] < 80:
            s += 0.2
        return s
    ls, rs = score(l, r), score(r, l)
    t = occlusion_threshold
    if ls > t and rs < t * 0.6:
        result = "left"
    elif rs > t and ls < t * 0.6:
        result = "right"
    elif ls > t and rs > t:
        result = "both"
    else:
        result = "none"
    return result, ls, rs


class OcclusionMonitor:
    """The occlusion hysteresis of the reference's main loop (fused_depth_map.py:2325-2337 state, :2515-2571 update):
    every `check_interval`-th frame the detector runs; a detected state that differs from the accepted one must
    be seen on `hysteresis_frames` consecutive checks before it is accepted (a check that agrees with the accepted
    state resets the count); accepting 'left' / 'right' switches stereo off, accepting 'none' / 'both' switches
    it back on (with both cameras covered the reference keeps stereo running).  Host logic on three scalars per
    check; the image statistics behind `detect` run on the device.

        mon = OcclusionMonitor()
        for frame_counter, (left, right) in enumerate(frames):
            use_stereo = mon.update(frame_counter, left, right)
    """

    def __init__(self, hysteresis_frames=5, check_interval=2, occlusion_threshold=0.45, use_stereo=True, detect=None):
        self.hysteresis_frames = int(hysteresis_frames)          # OCCLUSION_HYSTERESIS_FRAMES
        self.check_interval = int(check_interval)                # OCCLUSION_CHECK_INTERVAL
        self.occlusion_threshold = occlusion_threshold
        self.state = "none"                                      # occlusion_state
        self.counter = 0                                         # occlusion_hysteresis_counter
        self.use_stereo = bool(use_stereo)                       # USE_STEREO
        self.left_score = self.right_score = 0.0
        self.last_detected = "none"
        self._detect = detect or detect_camera_occlusion

    def update(self, frame_counter, left_img, right_img):
        """One iteration of the reference's loop body; returns whether stereo is enabled for this frame."""
        if frame_counter % self.check_interval != 0:
            return self.use_stereo
        new_state, self.left_score, self.right_score = self._detect(left_img, right_img,
                                                                    occlusion_threshold=self.occlusion_threshold)
        self.last_detected = new_state
        return self.observe(new_state)

    def observe(self, new_state):
        """The state machine alone, fed with a detector result."""
        if new_state == self.state:
            self.counter = 0
            return self.use_stereo
        self.counter += 1
        if self.counter >= self.hysteresis_frames:
            self.state = new_state
            self.counter = 0
            if self.state in ("left", "right"):
                self.use_stereo = False
            elif self.state in ("none", "both"):
                self.use_stereo = True
        return self.use_stereo
