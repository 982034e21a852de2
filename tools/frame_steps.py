#!/usr/bin/env python
"""Per-call times of the depth_map.py frame path (script parameters: D = 320, window 7, 3WAY), call by call with NumPy
arrays: which step is slow in the runs where the whole frame takes 6 instead of 3 ms."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import stereovision_b200 as svb
from stereovision_b200.synth import make_pair, to_bgr, synthetic_calibration
from stereovision_b200 import calib

H, W, D, bs = 480, 640, 320, 7
c = synthetic_calibration(W, H)
R1, R2, P1, P2, Q, _, _ = calib.stereoRectify(c["mtx_left"], c["dist_left"], c["mtx_right"], c["dist_right"], (W, H), c["R"], c["T"], alpha=0)
ml = svb.initUndistortRectifyMap(c["mtx_left"], c["dist_left"], R1, P1, (W, H))
mr = svb.initUndistortRectifyMap(c["mtx_right"], c["dist_right"], R2, P2, (W, H))
l, r, _ = make_pair(H, W, 128, seed=9)
lb, rb = to_bgr(l), to_bgr(r)
names = ["remap_l", "remap_r", "gray_l", "gray_r", "sgbm", "median", "reproject"]
acc = {n: [] for n in names}
for it in range(60):
    t = [time.perf_counter()]
    lr = svb.remap(lb, ml[0], ml[1], svb.INTER_LINEAR); t.append(time.perf_counter())
    rr = svb.remap(rb, mr[0], mr[1], svb.INTER_LINEAR); t.append(time.perf_counter())
    gl = svb.cvtColor(lr, svb.COLOR_BGR2GRAY); t.append(time.perf_counter())
    gr = svb.cvtColor(rr, svb.COLOR_BGR2GRAY); t.append(time.perf_counter())
    st = svb.StereoSGBM_create(minDisparity=0, numDisparities=D, blockSize=bs, P1=8 * 3 * bs * bs, P2=32 * 3 * bs * bs, disp12MaxDiff=1,
                               uniquenessRatio=10, speckleWindowSize=100, speckleRange=32, preFilterCap=63, mode=2)
    d16 = st.compute(gl, gr); t.append(time.perf_counter())
    disp = svb.medianBlur(d16.astype(np.float32) / 16.0, 5); t.append(time.perf_counter())
    pts = svb.reprojectImageTo3D(d16, Q); t.append(time.perf_counter())
    if it >= 30:
        for n, a, b in zip(names, t[:-1], t[1:]):
            acc[n].append((b - a) * 1e3)
print({n: round(sorted(v)[len(v) // 2], 3) for n, v in acc.items()}, "total", round(sum(sorted(v)[len(v) // 2] for v in acc.values()), 2))
