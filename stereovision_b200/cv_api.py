"""OpenCV-shaped surface of the engine: the calls the reference makes on its per-frame
path, with cv2's names, argument order and error behaviour, running on libsvb.so.

    cv2.StereoSGBM_create(...).compute   depth_map.py:894-909, fused_depth_map.py:988-1004
    cv2.StereoBM_create(...).compute     BASELINE.json config 2
    cv2.remap / cv2.cvtColor             depth_map.py:815-826, 871-879
    cv2.medianBlur / cv2.filterSpeckles  depth_map.py:912; inside StereoSGBM
    cv2.reprojectImageTo3D               BASELINE.json (Q from depth_map.py:625)

Inputs may be NumPy arrays (host: copied in and out, exactly what the reference passes)
or torch CUDA tensors (zero-copy, results stay on the device, work is enqueued on the
current torch stream).  Every function also accepts a leading batch dimension.
"""
import ctypes
import threading

import numpy as np

from . import _lib

STEREO_SGBM_MODE_SGBM = 0
STEREO_SGBM_MODE_HH = 1
STEREO_SGBM_MODE_SGBM_3WAY = 2
STEREO_SGBM_MODE_HH4 = 3
STEREO_BM_PREFILTER_NORMALIZED_RESPONSE, STEREO_BM_PREFILTER_XSOBEL = 0, 1
StereoSGBM_MODE_SGBM, StereoSGBM_MODE_HH = 0, 1
StereoSGBM_MODE_SGBM_3WAY, StereoSGBM_MODE_HH4 = 2, 3
INTER_LINEAR = 1
COLOR_BGR2GRAY = 6

_tls = threading.local()
_pool_lock = threading.Lock()
_pool = {}  # (thread, device, kind, params) -> handle


def _torch():
    import torch
    return torch


def _is_tensor(x):
    return type(x).__module__.startswith("torch")


def _ensure_device(device=None):
    """Make `device` (default: torch's current device) the CUDA device of the calling thread for the rest of the
    public call -- libsvb launches on the runtime's current device, and the reference calls from a pool thread
    (fused_depth_map.py:2591).  Compared against torch's live current device on every call, never a cached
    value; `_restores_device` puts the caller's device back when the public function returns."""
    torch = _torch()
    if not torch.cuda.is_available():
        raise _lib.SvbError("no CUDA device available; stereovision_b200 has no CPU fallback")
    cur = torch.cuda.current_device()
    if device is None:
        device = cur
    device = int(device)
    if device != cur:
        torch.cuda.set_device(device)
    seen = getattr(_tls, "inited", None)
    if seen is None:
        seen = _tls.inited = set()
    if device not in seen:                    # once per (thread, device): context + sm_100 check
        _lib.check(_lib.lib().svb_init(device))
        seen.add(device)
    return device


def _restores_device(fn):
    """Decorator of every public entry point: whatever device the call had to select (inputs living on another
    GPU than the thread's current one), the caller's current device is restored on the way out."""
    import functools

    @functools.wraps(fn)
    def wrapper(*args, **kwargs):
        torch = _torch()
        if not torch.cuda.is_available():
            return fn(*args, **kwargs)
        prev = torch.cuda.current_device()
        try:
            return fn(*args, **kwargs)
        finally:
            if torch.cuda.current_device() != prev:
                torch.cuda.set_device(prev)
    return wrapper


def _stream_ptr():
    torch = _torch()
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


_POOL_PER_THREAD = 4          # matcher handles (device workspaces) a thread keeps alive; least recently used goes first


def _destroy(kind, h):
    L = _lib.lib()
    (L.svb_sgbm_destroy if kind == "sgbm" else L.svb_bm_destroy)(h)


def _handle(kind, params, device):
    """The calling thread's handle for (device, kind, params).  The reference builds a new matcher object per frame
    (depth_map.py:894) and changes parameters live (fused_depth_map.py:3241-3289), so handles are pooled per thread
    (a handle is never shared between threads) and the pool is an LRU: a thread keeps at most _POOL_PER_THREAD
    handles, evicted ones -- and everything a finished thread left behind -- are destroyed (svb_*_destroy frees the
    device workspace)."""
    me = threading.get_ident()
    key = (me, device, kind, params)
    with _pool_lock:
        h = _pool.pop(key, None)
        if h is not None:
            _pool[key] = h                    # most recently used last (dicts keep insertion order)
            return h
    L = _lib.lib()
    out = ctypes.c_void_p()
    if kind == "sgbm":
        prm = _lib.SgbmParams(*params)
        _lib.check(L.svb_sgbm_create(ctypes.byref(prm), ctypes.byref(out)))
    else:
        prm = _lib.BmParams(*params)
        _lib.check(L.svb_bm_create(ctypes.byref(prm), ctypes.byref(out)))
    alive = {t.ident for t in threading.enumerate()}
    doomed = []
    with _pool_lock:
        _pool[key] = out
        mine = [k for k in _pool if k[0] == me]
        for k in mine[:max(0, len(mine) - _POOL_PER_THREAD)]:
            doomed.append((k[2], _pool.pop(k)))
        for k in [k for k in _pool if k[0] not in alive]:
            doomed.append((k[2], _pool.pop(k)))
    for kd, hd in doomed:                      # svb_*_destroy waits for the handle's own work (cudaFree)
        _destroy(kd, hd)
    return out


def release_workspaces():
    """Free every cached matcher handle (device workspaces).  Call it when no other thread is inside compute()."""
    with _pool_lock:
        items = list(_pool.items())
        _pool.clear()
    for (_, _, kind, _), h in items:
        _destroy(kind, h)


def pool_size():
    """Number of live matcher handles (all threads)."""
    with _pool_lock:
        return len(_pool)


def _check_pair(left, right):
    if _is_tensor(left) != _is_tensor(right):
        raise ValueError("left and right must both be NumPy arrays or both torch tensors")
    if tuple(left.shape) != tuple(right.shape):
        raise ValueError(f"left and right must have the same size: {tuple(left.shape)} vs {tuple(right.shape)}")
    if left.ndim not in (2, 3, 4) or (left.ndim == 4 and left.shape[-1] != 3):
        raise ValueError("expected [H,W] / [B,H,W] single-channel or [H,W,3] / [B,H,W,3] colour 8-bit images")
    if _is_tensor(left):
        torch = _torch()
        if left.dtype != torch.uint8 or right.dtype != torch.uint8:
            raise ValueError("images must be 8-bit (CV_8U)")
        if not left.is_cuda:
            raise ValueError("torch inputs must be CUDA tensors (use NumPy for host data)")
    else:
        if left.dtype != np.uint8 or right.dtype != np.uint8:
            raise ValueError("images must be 8-bit (CV_8U)")


def _row_pitches(left, right, cn):
    """(row_pitch, img_pitch) in bytes when both tensors are non-contiguous views whose pixels are contiguous within a
    row and whose strides agree (a crop of a larger frame, a padded buffer) -- what svb_sgbm_compute_pitched_dev
    takes without a copy; None for dense tensors or anything else (those go through .contiguous())."""
    if left.is_contiguous() and right.is_contiguous():
        return None
    if tuple(left.stride()) != tuple(right.stride()):
        return None
    st = list(left.stride())
    if cn == 3:
        if st[-1] != 1 or st[-2] != 3:
            return None
        st = st[:-1]
        st[-1] = 1
    if st[-1] != 1 or st[-2] < left.shape[-1 if cn == 1 else -2] * cn:
        return None
    row = int(st[-2])
    img = int(st[0]) if len(st) == 3 else 0
    if len(st) == 3 and left.shape[0] > 1 and img < row * (int(left.shape[1]) - 1) + int(left.shape[2]) * cn:
        return None
    return row, img


class _Matcher:
    _kind = None

    def _params(self):
        raise NotImplementedError

    @_restores_device
    def compute(self, left, right):
        """disparity = matcher.compute(left, right) -> int16 (x16 fixed point), same shape."""
        _check_pair(left, right)
        L = _lib.lib()
        # cv2 reads a 3-dim array as H x W x channels: a trailing dimension of 3 is colour (no image is 3 wide)
        cn = 3 if left.ndim >= 3 and left.shape[-1] == 3 else 1
        oshape = tuple(left.shape[:-1]) if cn == 3 else tuple(left.shape)
        B = oshape[0] if len(oshape) == 3 else 1
        H, W = oshape[-2], oshape[-1]
        if self._kind == "sgbm":
            host_fn = lambda h, l, r, B, H, W, o: L.svb_sgbm_compute_cn_host(h, l, r, B, H, W, cn, o)
            dev_fn = lambda h, l, r, B, H, W, o, st: L.svb_sgbm_compute_cn_dev(h, l, r, B, H, W, cn, o, st)
        else:
            if cn != 1:
                raise ValueError("StereoBM needs single-channel 8-bit images (cv2: CV_8UC1)")
            host_fn, dev_fn = L.svb_bm_compute_host, L.svb_bm_compute_dev
        if _is_tensor(left):
            torch = _torch()
            dev = _ensure_device(left.device.index)
            h = _handle(self._kind, self._params(), dev)
            self._pre_compute(h)
            out = torch.empty(oshape, dtype=torch.int16, device=left.device)
            pitches = _row_pitches(left, right, cn) if self._kind == "sgbm" else None
            if pitches is not None:                           # a crop / padded view: no copy, pitched C-ABI call
                _lib.check(L.svb_sgbm_compute_pitched_dev(h, left.data_ptr(), right.data_ptr(), B, H, W, cn, pitches[0],
                                                          pitches[1], out.data_ptr(), _stream_ptr()))
            else:
                l, r = left.contiguous(), right.contiguous()
                _lib.check(dev_fn(h, l.data_ptr(), r.data_ptr(), B, H, W, out.data_ptr(), _stream_ptr()))
            self._last = h
            return out
        dev = _ensure_device()
        h = _handle(self._kind, self._params(), dev)
        self._pre_compute(h)
        l = np.ascontiguousarray(left)
        r = np.ascontiguousarray(right)
        out = np.empty(oshape, np.int16)
        _lib.check(host_fn(h, l.ctypes.data, r.ctypes.data, B, H, W, out.ctypes.data))
        self._last = h
        return out

    def _pre_compute(self, h):
        pass

    def check(self):
        """Wait for this thread's last compute() with these parameters and raise if the aggregation kernels flagged
        a timed-out inter-CTA wait (svb_sgbm_check).  NumPy calls check before they return; for torch-tensor
        calls (asynchronous, no hidden sync) this is the caller's hook at its own synchronisation point."""
        h = getattr(self, "_last", None)
        if h is not None and self._kind == "sgbm":
            _lib.check(_lib.lib().svb_sgbm_check(h))

    def last_launches(self):
        L = _lib.lib()
        h = getattr(self, "_last", None)
        if h is None:
            return 0
        return (L.svb_sgbm_last_launches if self._kind == "sgbm" else L.svb_bm_last_launches)(h)


class StereoSGBM(_Matcher):
    """Parameter holder, cheap to construct every frame as the reference does
    (depth_map.py:894); device workspaces live in a module-level pool."""
    _kind = "sgbm"
    _names = ("minDisparity", "numDisparities", "blockSize", "P1", "P2", "disp12MaxDiff",
              "preFilterCap", "uniquenessRatio", "speckleWindowSize", "speckleRange", "mode")

    def __init__(self, minDisparity=0, numDisparities=16, blockSize=3, P1=0, P2=0, disp12MaxDiff=0,
                 preFilterCap=0, uniquenessRatio=0, speckleWindowSize=0, speckleRange=0,
                 mode=STEREO_SGBM_MODE_SGBM):
        for n in self._names:
            setattr(self, "_" + n, int(locals()[n]))

    def _params(self):
        return tuple(getattr(self, "_" + n) for n in self._names)

    def setDebugStages(self, enable=True):
        """Test hook: keep the aggregated volume S of the next compute() in device memory
        (normally the last sweep consumes it in registers) so debug_stages() can return it."""
        self._debug = bool(enable)

    def setDebugForceWave(self, mode):
        """Test hook: 1 / 0 = the rows-per-warp / warp-per-row kernel for the first sweep at numDisparities <= 128
        whatever the batch size, -1 = automatic (svb_sgbm_debug_force_wave)."""
        self._force_wave = int(mode)

    def _pre_compute(self, h):
        _lib.check(_lib.lib().svb_sgbm_set_debug(h, int(getattr(self, "_debug", False))))
        _lib.check(_lib.lib().svb_sgbm_debug_force_wave(h, int(getattr(self, "_force_wave", -1))))

    def debug_stages(self, H, W):
        """Test hook: (C, S, raw) of image 0 of the last compute as NumPy arrays
        (call setDebugStages() before compute())."""
        L = _lib.lib()
        D, minD = self._numDisparities, self._minDisparity
        w1 = W + min(minD, 0) - max(minD + D, 0)
        C = np.zeros((H, w1, D), np.uint16)
        S = np.zeros((H, w1, D), np.uint16)
        raw = np.zeros((H, W), np.int16)
        _lib.check(L.svb_sgbm_debug_stages(self._last, C.ctypes.data, S.ctypes.data, raw.ctypes.data))
        return C, S, raw


def _add_accessors(cls, names):
    for n in names:
        cap = n[0].upper() + n[1:]
        setattr(cls, "get" + cap, (lambda n: lambda self: getattr(self, "_" + n))(n))
        setattr(cls, "set" + cap, (lambda n: lambda self, v: setattr(self, "_" + n, int(v)))(n))


_add_accessors(StereoSGBM, StereoSGBM._names)


def StereoSGBM_create(minDisparity=0, numDisparities=16, blockSize=3, P1=0, P2=0, disp12MaxDiff=0,
                      preFilterCap=0, uniquenessRatio=0, speckleWindowSize=0, speckleRange=0,
                      mode=STEREO_SGBM_MODE_SGBM):
    """cv2.StereoSGBM_create with cv2's positional order and defaults."""
    return StereoSGBM(minDisparity, numDisparities, blockSize, P1, P2, disp12MaxDiff, preFilterCap,
                      uniquenessRatio, speckleWindowSize, speckleRange, mode)


class StereoBM(_Matcher):
    """cv2.StereoBM_create(numDisparities=0 -> 64, blockSize=21) with cv2's defaults
    (preFilterCap 31, textureThreshold 10, uniquenessRatio 15, X-Sobel prefilter)."""
    _kind = "bm"
    _names = ("minDisparity", "numDisparities", "blockSize", "preFilterCap", "textureThreshold",
              "uniquenessRatio", "speckleWindowSize", "speckleRange", "disp12MaxDiff", "preFilterType",
              "preFilterSize")

    def __init__(self, numDisparities=0, blockSize=21):
        self._minDisparity = 0
        self._numDisparities = int(numDisparities) if numDisparities > 0 else 64
        self._blockSize = int(blockSize)
        self._preFilterCap = 31
        self._textureThreshold = 10
        self._uniquenessRatio = 15
        self._speckleWindowSize = 0
        self._speckleRange = 0
        self._disp12MaxDiff = -1
        self._preFilterType = STEREO_BM_PREFILTER_XSOBEL
        self._preFilterSize = 9

    def _params(self):
        return tuple(getattr(self, "_" + n) for n in self._names)


_add_accessors(StereoBM, StereoBM._names)


def StereoBM_create(numDisparities=0, blockSize=21):
    return StereoBM(numDisparities, blockSize)


# --------------------------------------------------------------------------------------
# imgproc / calib3d helpers
# --------------------------------------------------------------------------------------
def _to_dev(a, dtype=None):
    """NumPy -> CUDA tensor (plumbing only); tensors pass through."""
    torch = _torch()
    if _is_tensor(a):
        if not a.is_cuda:
            raise ValueError("torch inputs must be CUDA tensors")
        return a.contiguous(), True
    if not torch.cuda.is_available():
        raise _lib.SvbError("no CUDA device available; stereovision_b200 has no CPU fallback")
    a = np.ascontiguousarray(a if dtype is None else a.astype(dtype, copy=False))
    return _upload(a), False


# NumPy arrays above this size travel through svb_copy_h2d / svb_copy_d2h (pinned staging filled by a few host threads
# while the previous chunk is on the bus): a 1080p BGR batch arrives ~3x faster than through a pageable cudaMemcpy.
_STAGED_COPY_BYTES = 4 << 20


def _upload(a):
    """Contiguous NumPy array -> CUDA tensor on the current device and stream."""
    torch = _torch()
    if a.nbytes < _STAGED_COPY_BYTES or not a.flags.c_contiguous:
        return torch.from_numpy(a).cuda(non_blocking=False)
    _ensure_device()
    t = torch.empty(a.shape, dtype=torch.from_numpy(a[:0].reshape(-1)).dtype, device="cuda")
    _lib.check(_lib.lib().svb_copy_h2d(t.data_ptr(), a.ctypes.data, a.nbytes, _stream_ptr()))
    return t


def _download(t):
    """CUDA tensor -> new NumPy array (waits for the tensor's producer on the current stream)."""
    torch = _torch()
    n = t.numel() * t.element_size()
    if n < _STAGED_COPY_BYTES:
        return t.cpu().numpy()
    t = t.contiguous()
    out = np.empty(tuple(t.shape), torch.empty(0, dtype=t.dtype).numpy().dtype)
    with torch.cuda.device(t.device):
        _lib.check(_lib.lib().svb_copy_d2h(out.ctypes.data, t.data_ptr(), n, _stream_ptr()))
    return out


def _ret(t, was_tensor):
    return t if was_tensor else _download(t)


# Rectification maps are built once per calibration (depth_map.py:633-638) and passed to cv2.remap on every
# frame: keep their device copies.  Keyed by the array object (dropped with it through a weakref); a strided
# sample of the contents guards against in-place edits.
_map_cache = {}
_map_cache_lock = threading.Lock()


_MAP_FULL_CHECK_BYTES = 4 << 20


def _map_sample(a):
    """Fingerprint of a map's contents: the whole array for maps up to 4 MB (a 640x480 CV_16SC2 map is 1.2 MB,
    about 0.1 ms), XOR-folded as 64-bit words; 4096 strided elements plus the tail beyond that (hashing 12 MB of
    1080p maps on every frame would cost more than the remap).  In-place recalibration of a large map should go
    through invalidate_map_cache() -- or keep the maps on the device (initUndistortRectifyMap(as_tensor=True),
    load_stereo_calibration(device_maps=True)), which bypasses the cache."""
    flat = a.reshape(-1)
    if flat.nbytes <= _MAP_FULL_CHECK_BYTES:
        raw = flat.view(np.uint8)
        n8 = raw.size // 8 * 8
        fold = int(np.bitwise_xor.reduce(raw[:n8].view(np.uint64))) if n8 else 0
        return (fold, int(raw[:n8].view(np.uint64)[::7].sum(dtype=np.uint64)) if n8 else 0, raw[n8:].tobytes())
    step = max(1, flat.size // 4096)
    return flat[::step].tobytes() + flat[-64:].tobytes()


def invalidate_map_cache():
    """Drop every cached device copy of a rectification map."""
    with _map_cache_lock:
        _map_cache.clear()


def _cached_map(a):
    """Device copy of a NumPy map (cached); tensors pass through."""
    if _is_tensor(a) or not isinstance(a, np.ndarray) or not a.flags.c_contiguous:
        return _to_dev(a)[0]
    import weakref
    torch = _torch()
    dev = torch.cuda.current_device() if torch.cuda.is_available() else -1
    key = (id(a), dev)
    sample = _map_sample(a)
    with _map_cache_lock:
        hit = _map_cache.get(key)
    if hit is not None and hit[1] == (a.shape, a.dtype.str, a.ctypes.data) and hit[2] == sample:
        return hit[0]
    t = _to_dev(a)[0]
    try:
        ref = weakref.ref(a, lambda _r, k=key: _map_cache.pop(k, None))
    except TypeError:
        return t
    with _map_cache_lock:
        if len(_map_cache) > 64:
            _map_cache.clear()
        _map_cache[key] = (t, (a.shape, a.dtype.str, a.ctypes.data), sample, ref)
    return t


@_restores_device
def remap(src, map1, map2, interpolation=INTER_LINEAR, to_gray=False):
    """cv2.remap(src, map1, map2, cv2.INTER_LINEAR) for the fixed-point maps the reference
    builds with initUndistortRectifyMap(..., CV_16SC2) (depth_map.py:633-638, 815-826).
    src: u8 [H,W], [H,W,3], or batched [B,H,W(,3)].  to_gray=True additionally applies
    cv2.cvtColor(BGR2GRAY) to the remapped image (depth_map.py:871-879) in the same kernel."""
    if interpolation != INTER_LINEAR:
        raise NotImplementedError("only INTER_LINEAR is on the reference's path")
    torch = _torch()
    m1 = _cached_map(map1)
    m2 = _cached_map(map2)
    if m1.dtype != torch.int16 or m1.ndim != 3 or m1.shape[2] != 2 or m2.dtype not in (torch.uint16, torch.int16):
        raise ValueError("maps must be CV_16SC2 [H,W,2] int16 + CV_16UC1 [H,W] uint16")
    s, was = _to_dev(src)
    if s.dtype != torch.uint8:
        raise ValueError("src must be 8-bit")
    Ho, Wo = int(m2.shape[0]), int(m2.shape[1])
    # layouts: [H,W] [H,W,3] [B,H,W] [B,H,W,3]; a trailing dim of 3 means channels
    if s.ndim == 2:
        B, C, batched = 1, 1, False
    elif s.ndim == 3 and s.shape[-1] == 3:
        B, C, batched = 1, 3, False
    elif s.ndim == 3:
        B, C, batched = int(s.shape[0]), 1, True
    elif s.ndim == 4 and s.shape[-1] == 3:
        B, C, batched = int(s.shape[0]), 3, True
    else:
        raise ValueError(f"unsupported src shape {tuple(s.shape)}")
    Hs, Ws = (int(s.shape[1]), int(s.shape[2])) if batched else (int(s.shape[0]), int(s.shape[1]))
    dev = _ensure_device(s.device.index)
    oshape = ([B] if batched else []) + [Ho, Wo] + ([3] if (C == 3 and not to_gray) else [])
    out = torch.empty(oshape, dtype=torch.uint8, device=s.device)
    _lib.check(_lib.lib().svb_remap_u8_dev(s.data_ptr(), B, Hs, Ws, C, m1.data_ptr(), m2.data_ptr(), Ho, Wo,
                                           out.data_ptr(), int(bool(to_gray)), _stream_ptr()))
    return _ret(out, was)


@_restores_device
def resize(src, dsize, dst=None, fx=0, fy=0, interpolation=INTER_LINEAR):
    """cv2.resize(src, dsize, interpolation=cv2.INTER_LINEAR) -- cv2's positional order; INTER_LINEAR is also
    cv2's default (depth_map.py:776).  dsize = (width, height); dsize None/(0, 0) takes round(fx * W), round(fy * H)
    like cv2.  u8 [H,W] / [H,W,3] / [B,H,W] / [B,H,W,3] (bit-exact; fused_depth_map.py:473-476, depth_map.py:809-811)
    or float32 [H,W] / [B,H,W] (fused_depth_map.py:1214-1217).  A source that already has the size comes back as a
    copy, as from cv2."""
    if interpolation != INTER_LINEAR:
        raise NotImplementedError("only INTER_LINEAR is on the reference's path")
    if dst is not None:
        raise NotImplementedError("resize: pass no dst (the result is returned)")
    torch = _torch()
    s, was = _to_dev(src)
    if s.dtype not in (torch.uint8, torch.float32):
        raise ValueError("resize: src must be 8-bit or float32")
    if s.ndim == 2:
        B, C, batched = 1, 1, False
    elif s.ndim == 3 and s.shape[-1] == 3 and s.dtype == torch.uint8:
        B, C, batched = 1, 3, False
    elif s.ndim == 3:
        B, C, batched = int(s.shape[0]), 1, True
    elif s.ndim == 4 and s.shape[-1] == 3 and s.dtype == torch.uint8:
        B, C, batched = int(s.shape[0]), 3, True
    else:
        raise ValueError(f"unsupported src shape {tuple(s.shape)}")
    Hs, Ws = (int(s.shape[1]), int(s.shape[2])) if batched else (int(s.shape[0]), int(s.shape[1]))
    if dsize is None or tuple(dsize) == (0, 0):
        if not (fx > 0 and fy > 0):
            raise ValueError("resize: either dsize or both fx and fy must be non-zero (cv2: inv_scale_x > 0)")
        Wd, Hd = int(round(Ws * fx)), int(round(Hs * fy))     # cv2: saturate_cast<int>(ssize * inv_scale)
    else:
        Wd, Hd = int(dsize[0]), int(dsize[1])
    _ensure_device(s.device.index)
    oshape = ([B] if batched else []) + [Hd, Wd] + ([3] if C == 3 else [])
    out = torch.empty(oshape, dtype=s.dtype, device=s.device)
    L = _lib.lib()
    if s.dtype == torch.uint8:
        _lib.check(L.svb_resize_u8_dev(s.data_ptr(), B, Hs, Ws, C, out.data_ptr(), Hd, Wd, _stream_ptr()))
    else:
        _lib.check(L.svb_resize_f32_dev(s.data_ptr(), B, Hs, Ws, out.data_ptr(), Hd, Wd, _stream_ptr()))
    return _ret(out, was)


@_restores_device
def cvtColor(src, code=COLOR_BGR2GRAY):
    """cv2.cvtColor(img, cv2.COLOR_BGR2GRAY) (depth_map.py:871-879)."""
    if code != COLOR_BGR2GRAY:
        raise NotImplementedError("only COLOR_BGR2GRAY is on the reference's path")
    torch = _torch()
    s, was = _to_dev(src)
    if s.dtype != torch.uint8 or s.shape[-1] != 3:
        raise ValueError("expected 8-bit BGR [...,3]")
    _ensure_device(s.device.index)
    out = torch.empty(s.shape[:-1], dtype=torch.uint8, device=s.device)
    _lib.check(_lib.lib().svb_bgr2gray_dev(s.data_ptr(), out.numel(), out.data_ptr(), _stream_ptr()))
    return _ret(out, was)


@_restores_device
def medianBlur(src, ksize):
    """cv2.medianBlur: ksize 3 on int16 (as inside StereoSGBM) or 5 on float32 (depth_map.py:912)."""
    torch = _torch()
    s, was = _to_dev(src)
    if s.ndim not in (2, 3):
        raise ValueError("expected [H,W] or [B,H,W]")
    B = 1 if s.ndim == 2 else int(s.shape[0])
    H, W = int(s.shape[-2]), int(s.shape[-1])
    _ensure_device(s.device.index)
    out = torch.empty_like(s)
    L = _lib.lib()
    if ksize == 3 and s.dtype == torch.int16:
        _lib.check(L.svb_median3_s16_dev(s.data_ptr(), B, H, W, out.data_ptr(), _stream_ptr()))
    elif ksize == 5 and s.dtype == torch.float32:
        _lib.check(L.svb_median5_f32_dev(s.data_ptr(), B, H, W, out.data_ptr(), _stream_ptr()))
    else:
        raise NotImplementedError("medianBlur: only (ksize=3, int16) and (ksize=5, float32) are on the path")
    return _ret(out, was)


@_restores_device
def filterSpeckles(img, newVal, maxSpeckleSize, maxDiff):
    """cv2.filterSpeckles(img, newVal, maxSpeckleSize, maxDiff) on int16; returns the filtered
    image (tensors are modified in place like cv2 does, NumPy inputs get a filtered copy back
    AND are updated in place to keep cv2's in-place contract)."""
    torch = _torch()
    s, was = _to_dev(img)
    if s.dtype != torch.int16 or s.ndim not in (2, 3):
        raise ValueError("expected int16 [H,W] or [B,H,W]")
    B = 1 if s.ndim == 2 else int(s.shape[0])
    H, W = int(s.shape[-2]), int(s.shape[-1])
    _ensure_device(s.device.index)
    scratch = torch.empty(2 * s.numel(), dtype=torch.int32, device=s.device)
    _lib.check(_lib.lib().svb_filter_speckles_s16_dev(s.data_ptr(), B, H, W, int(newVal), int(maxSpeckleSize),
                                                      int(maxDiff), scratch.data_ptr(), _stream_ptr()))
    if was:
        return s
    out = s.cpu().numpy()
    img[...] = out
    return img


@_restores_device
def validateDisparity(disparity, cost, minDisparity, numberOfDisparities, disp12MaxDiff=1):
    """cv2.validateDisparity(disparity, cost, minDisparity, numberOfDisparities[, disp12MaxDiff]): the
    left-right check StereoBM runs when disp12MaxDiff >= 0.  disparity int16 (x16), cost int16 or int32,
    [H,W] or [B,H,W]; modified in place like cv2 and returned."""
    torch = _torch()
    d, was = _to_dev(disparity)
    c, _ = _to_dev(cost)
    if d.dtype != torch.int16 or d.ndim not in (2, 3):
        raise ValueError("disparity: expected int16 [H,W] or [B,H,W]")
    if c.dtype not in (torch.int16, torch.int32) or tuple(c.shape) != tuple(d.shape):
        raise ValueError("cost: expected int16 or int32 of the disparity's shape")
    B = 1 if d.ndim == 2 else int(d.shape[0])
    H, W = int(d.shape[-2]), int(d.shape[-1])
    _ensure_device(d.device.index)
    scratch = torch.empty(10 * d.numel() + 16, dtype=torch.uint8, device=d.device)
    _lib.check(_lib.lib().svb_validate_disparity_dev(d.data_ptr(), c.data_ptr(), int(c.dtype == torch.int32), B, H, W,
                                                     int(minDisparity), int(numberOfDisparities), int(disp12MaxDiff),
                                                     scratch.data_ptr(), _stream_ptr()))
    if was:
        return d
    disparity[...] = d.cpu().numpy()
    return disparity


@_restores_device
def reprojectImageTo3D(disparity, Q, handleMissingValues=False):
    """cv2.reprojectImageTo3D(disparity, Q): int16 input is NOT divided by 16."""
    torch = _torch()
    d, was = _to_dev(disparity)
    if d.dtype not in (torch.int16, torch.float32) or d.ndim not in (2, 3):
        raise ValueError("disparity must be int16 or float32, [H,W] or [B,H,W]")
    Qh = np.ascontiguousarray(Q.detach().cpu().numpy() if _is_tensor(Q) else Q, dtype=np.float64)
    if Qh.shape != (4, 4):
        raise ValueError("Q must be 4x4")
    B = 1 if d.ndim == 2 else int(d.shape[0])
    H, W = int(d.shape[-2]), int(d.shape[-1])
    _ensure_device(d.device.index)
    out = torch.empty(tuple(d.shape) + (3,), dtype=torch.float32, device=d.device)
    scratch = torch.empty(B, dtype=torch.float32, device=d.device)
    _lib.check(_lib.lib().svb_reproject3d_dev(
        d.data_ptr(), int(d.dtype == torch.int16), B, H, W, Qh.ctypes.data_as(ctypes.POINTER(ctypes.c_double)),
        int(bool(handleMissingValues)), scratch.data_ptr(), out.data_ptr(), _stream_ptr()))
    return _ret(out, was)


@_restores_device
def disparity_to_depth(disp16, fx=700.0, baseline=0.08, min_disp=0.0, min_depth=0.3, max_depth=2.0):
    """The tail of depth_map.create_depth_map (depth_map.py:909-933) in one kernel:
    disparity = disp16/16 -> medianBlur 5 -> depth = fx*B/(d+1e-6), clip, validity mask.
    Returns (depth_final, disparity f32).  depth_final is float32 for a Python-number fx and float64 for a NumPy
    float64 fx (camera_matrix[0, 0], depth_map.py:917) -- NumPy's promotion in the reference."""
    torch = _torch()
    f64 = isinstance(fx, np.floating) and not isinstance(fx, (np.float32, np.float16))
    d, was = _to_dev(disp16)
    if d.dtype != torch.int16 or d.ndim not in (2, 3):
        raise ValueError("disp16 must be int16 [H,W] or [B,H,W]")
    B = 1 if d.ndim == 2 else int(d.shape[0])
    H, W = int(d.shape[-2]), int(d.shape[-1])
    _ensure_device(d.device.index)
    disparity = torch.empty(d.shape, dtype=torch.float32, device=d.device)
    depth = torch.empty(d.shape, dtype=torch.float64 if f64 else torch.float32, device=d.device)
    fn = _lib.lib().svb_disp_to_depth_f64_dev if f64 else _lib.lib().svb_disp_to_depth_dev
    _lib.check(fn(d.data_ptr(), B, H, W, float(fx), float(baseline), float(min_disp),
                  float(min_depth), float(max_depth), disparity.data_ptr(), depth.data_ptr(), _stream_ptr()))
    return _ret(depth, was), _ret(disparity, was)


CV_16SC2 = 11


@_restores_device
def initUndistortRectifyMap(cameraMatrix, distCoeffs, R, newCameraMatrix, size, m1type=CV_16SC2, device=None,
                            as_tensor=False):
    """cv2.initUndistortRectifyMap(K, dist, R, P, (W, H), cv2.CV_16SC2) -> (map1 s16 [H,W,2], map2 u16 [H,W])
    (depth_map.py:633-638, fused_depth_map.py:404-409).  Computed on the device in float64 like OpenCV;
    as_tensor=True leaves the maps on the device for svb.remap."""
    if m1type != CV_16SC2:
        raise NotImplementedError("only CV_16SC2 maps are on the reference's path")
    torch = _torch()
    K = np.ascontiguousarray(cameraMatrix, dtype=np.float64)
    P = np.ascontiguousarray(newCameraMatrix, dtype=np.float64)
    if K.shape != (3, 3) or P.ndim != 2 or P.shape[0] != 3 or P.shape[1] not in (3, 4):
        raise ValueError("cameraMatrix must be 3x3 and newCameraMatrix 3x3 or 3x4")
    D = None if distCoeffs is None else np.ascontiguousarray(distCoeffs, dtype=np.float64).ravel()
    Rm = None if R is None else np.ascontiguousarray(R, dtype=np.float64)
    if Rm is not None and Rm.shape != (3, 3):
        raise ValueError("R must be 3x3")
    W, H = int(size[0]), int(size[1])
    if not torch.cuda.is_available():
        raise _lib.SvbError("no CUDA device available; stereovision_b200 has no CPU fallback")
    dev = _ensure_device(device)
    tdev = torch.device("cuda", dev)
    m1 = torch.empty((H, W, 2), dtype=torch.int16, device=tdev)
    m2 = torch.empty((H, W), dtype=torch.uint16, device=tdev)
    dp = ctypes.POINTER(ctypes.c_double)
    _lib.check(_lib.lib().svb_init_undistort_rectify_map_dev(
        K.ctypes.data_as(dp), D.ctypes.data_as(dp) if D is not None and D.size else None, 0 if D is None else int(D.size),
        Rm.ctypes.data_as(dp) if Rm is not None else None, P.ctypes.data_as(dp), int(P.shape[1]), W, H,
        m1.data_ptr(), m2.data_ptr(), _stream_ptr()))
    if as_tensor:
        return m1, m2
    return m1.cpu().numpy(), m2.cpu().numpy()
