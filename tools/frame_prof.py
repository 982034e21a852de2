import sys, os, time, numpy as np
sys.path.insert(0, "/root/repo")
import stereovision_b200 as svb
from stereovision_b200.synth import make_pair, to_bgr
from tests.golden_cases import synth_calibration
H, W = 480, 640
cal = synth_calibration(W, H)
l, r, _ = make_pair(H, W, 64, seed=9); lb, rb = to_bgr(l), to_bgr(r)
def t(name, fn, n=50):
    fn(); fn()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    print(f"{name:28s} {(time.perf_counter()-t0)/n*1e3:7.3f} ms")
t("remap bgr", lambda: svb.remap(lb, cal["left_map1"], cal["left_map2"], svb.INTER_LINEAR))
lr = svb.remap(lb, cal["left_map1"], cal["left_map2"], svb.INTER_LINEAR)
t("cvtColor", lambda: svb.cvtColor(lr, svb.COLOR_BGR2GRAY))
for D, bs, mode in ((320, 7, 2), (64, 5, 2), (64, 5, 0)):
    st = lambda: svb.StereoSGBM_create(minDisparity=0, numDisparities=D, blockSize=bs, P1=8*3*bs*bs, P2=32*3*bs*bs, disp12MaxDiff=1, uniquenessRatio=10, speckleWindowSize=100, speckleRange=32, preFilterCap=63, mode=mode).compute(l, r)
    t(f"sgbm D={D} bs={bs} mode={mode}", st)
d16 = svb.StereoSGBM_create(0, 64, 5).compute(l, r)
f = d16.astype(np.float32) / 16
t("astype/16 (numpy)", lambda: d16.astype(np.float32) / 16)
t("medianBlur 5", lambda: svb.medianBlur(f, 5))
t("reprojectImageTo3D", lambda: svb.reprojectImageTo3D(d16, cal["Q"]))
import torch
x = torch.from_numpy(lb)
t("torch H2D 0.9MB pageable", lambda: x.cuda())
y = x.cuda()
t("torch D2H 0.9MB", lambda: y.cpu())
