#!/usr/bin/env python
"""BASELINE.json configs[0]: one live frame through the depth_map.py path (640x480 BGR pair, synthetic calibration):
calibration .pkl on disk -> load_stereo_calibration (host stereoRectify + device maps), then per frame
remap x2 -> BGR2GRAY x2 -> StereoSGBM -> /16 -> medianBlur 5 -> depth, and reprojectImageTo3D.
Host arrays in, host arrays out, one frame at a time: the latency the reference's camera loop sees.  Times the
engine (cv2-shaped Python surface, the INTEGRATION.md binding) and cv2 itself on the same frame; checks the
int16 disparity bit for bit.  Not the headline bench (bench.py is)."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import stereovision_b200 as svb
from stereovision_b200.synth import make_pair, to_bgr, write_calibration_pkl

H, W = 480, 640
CASES = {"script": dict(D=320, bs=7, mode=2), "c1": dict(D=64, bs=5, mode=2), "c1_sgbm": dict(D=64, bs=5, mode=0)}


def frame(api, lb, rb, cal, D, bs, mode):
    lr = api.remap(lb, cal["left_map1"], cal["left_map2"], api.INTER_LINEAR)
    rr = api.remap(rb, cal["right_map1"], cal["right_map2"], api.INTER_LINEAR)
    gl, gr = api.cvtColor(lr, api.COLOR_BGR2GRAY), api.cvtColor(rr, api.COLOR_BGR2GRAY)
    st = api.StereoSGBM_create(minDisparity=0, numDisparities=D, blockSize=bs, P1=8 * 3 * bs * bs, P2=32 * 3 * bs * bs,
                               disp12MaxDiff=1, uniquenessRatio=10, speckleWindowSize=100, speckleRange=32,
                               preFilterCap=63, mode=mode)
    d16 = st.compute(gl, gr)
    disp = api.medianBlur(d16.astype(np.float32) / 16.0, 5)
    depth = (700 * 0.08) / (disp + 1e-6)
    pts = api.reprojectImageTo3D(d16, cal["Q"])
    return d16, depth, pts


def timeit(fn, n, warm=1):
    """Median of n single-call times after `warm` untimed calls (the first case otherwise runs while the GPU clocks
    still ramp up from idle: its mean used to swing between 3 and 7 ms from run to run)."""
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        fn()
        ts.append((time.perf_counter() - t0) * 1e3)
    return sorted(ts)[len(ts) // 2]


def main():
    import cv2
    import tempfile
    with tempfile.TemporaryDirectory() as tmp:               # the reference's file: output/stereo_calibration_data.pkl
        path = os.path.join(tmp, "output", "stereo_calibration_data.pkl")
        write_calibration_pkl(path, W, H)
        svb.load_stereo_calibration(path)                    # warm-up (context, kernels)
        t0 = time.perf_counter()
        cal = svb.load_stereo_calibration(path)
        load_ms = (time.perf_counter() - t0) * 1e3
    import torch
    dcal = {k: (torch.from_numpy(v).cuda() if isinstance(v, np.ndarray) and k.endswith(("map1", "map2")) else v) for k, v in cal.items()}
    out = {}
    # One throw-away case, engine and cv2: whichever case runs FIRST in a fresh process is ~3 ms slower per frame call
    # by call (same parameters 5.8 vs 3.0 ms when moved to second place) -- every NumPy result of a frame is a fresh
    # multi-megabyte allocation that glibc serves with mmap + page faults until its dynamic mmap threshold has adapted,
    # which the first large frees (cv2's included) bring about.  The reference's loop lives in the adapted state.
    l0, r0, _ = make_pair(H, W, 64, seed=8)
    for _ in range(5):
        frame(svb, to_bgr(l0), to_bgr(r0), cal, **CASES["c1"])
    frame(cv2, to_bgr(l0), to_bgr(r0), cal, **CASES["c1"])
    for name, c in CASES.items():
        l, r, _ = make_pair(H, W, min(c["D"], 128), seed=9)
        lb, rb = to_bgr(l), to_bgr(r)
        a = frame(svb, lb, rb, cal, **c)
        def frame_api():                                     # svb.frame: one upload, one download (device-resident maps)
            lr, rr = svb.frame.apply_stereo_rectification(lb, rb, dcal, as_tensor=True)
            return svb.frame.create_depth_map(lr, rr, None, min_disp=0, num_disp=c["D"], window_size=c["bs"])
        # the engine's legs first: cv2's worker threads keep spinning for a while after a call and share the host cores
        t_svb = timeit(lambda: frame(svb, lb, rb, cal, **c), 30, warm=30)
        t_api = timeit(frame_api, 30, warm=30) if c["mode"] == 2 else None
        b = frame(cv2, lb, rb, cal, **c)
        assert np.array_equal(a[0], b[0]), name
        fin = np.isfinite(b[2])
        assert np.allclose(a[2][fin], b[2][fin], rtol=1e-4, atol=1e-5)
        out[name] = {"svb_ms_per_frame": t_svb, "svb_frame_api_ms_per_frame": t_api,
                     "cv2_ms_per_frame": timeit(lambda: frame(cv2, lb, rb, cal, **c), 3), **c}
        time.sleep(0.5)
    print(json.dumps({"workload": "depth_map.py frame path 640x480 (BASELINE configs[0]), calibration from a .pkl",
                      "load_stereo_calibration_ms": load_ms, "cv2_threads": cv2.getNumThreads(),
                      "frames": out}))


if __name__ == "__main__":
    main()
