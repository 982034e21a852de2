import sys, os
sys.path.insert(0, "/root/repo")
import numpy as np, torch
import stereovision_b200 as svb
from stereovision_b200.synth import make_pair
for H, W, D, mode in [(480, 640, 64, 2), (480, 640, 64, 0), (480, 640, 128, 0), (720, 1280, 128, 0), (1080, 1920, 64, 2)]:
    l, r, _ = make_pair(H, W, D, seed=1)
    L, R = torch.from_numpy(l).cuda(), torch.from_numpy(r).cuda()
    m = svb.StereoSGBM_create(minDisparity=0, numDisparities=D, blockSize=5, P1=200, P2=800, disp12MaxDiff=1, preFilterCap=63,
                              uniquenessRatio=10, speckleWindowSize=100, speckleRange=32, mode=mode)
    for _ in range(3): m.compute(L, R)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(20):
        e0.record(); m.compute(L, R); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    print(f"{H}x{W} D{D} mode{mode} single frame: median {sorted(ts)[10]:.3f} ms")
