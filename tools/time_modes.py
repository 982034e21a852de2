#!/usr/bin/env python
"""Device-path time and output hash of StereoSGBM for a few (size, numDisparities, mode) combinations; used with
SVB_LIBRARY=build/libsvb_<tag>.so to A/B a kernel variant on configurations bench.py has no workload for.
   python tools/time_modes.py [--iters 5]"""
import argparse
import hashlib
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import stereovision_b200 as svb
from stereovision_b200.synth import make_pair

ap = argparse.ArgumentParser()
ap.add_argument("--iters", type=int, default=5)
args = ap.parse_args()
CASES = [(720, 1280, 128, 1, 32), (720, 1280, 128, 2, 32), (480, 640, 64, 1, 64), (480, 640, 64, 2, 64), (480, 640, 64, 3, 64),
         (480, 640, 96, 0, 64), (1080, 1920, 64, 1, 16)]
MODE = {0: "SGBM", 1: "HH", 2: "3WAY", 3: "HH4"}
for H, W, D, mode, B in CASES:
    pairs = [make_pair(H, W, D, seed=s) for s in range(4)]
    L = torch.from_numpy(np.stack([pairs[i % 4][0] for i in range(B)])).cuda()
    R = torch.from_numpy(np.stack([pairs[i % 4][1] for i in range(B)])).cuda()
    m = svb.StereoSGBM_create(minDisparity=0, numDisparities=D, blockSize=5, P1=200, P2=800, disp12MaxDiff=1, preFilterCap=63,
                              uniquenessRatio=10, speckleWindowSize=100, speckleRange=32, mode=mode)
    out = m.compute(L, R)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.iters):
        out = m.compute(L, R)
    e1.record()
    torch.cuda.synchronize()
    m.check()
    ms = e0.elapsed_time(e1) / args.iters
    h = hashlib.sha256(out.cpu().numpy().tobytes()).hexdigest()[:12]
    print(f"{H}x{W} D{D} {MODE[mode]:5s} B{B}: {ms:8.3f} ms  {B / ms * 1e3:9.1f} pairs/s  sha {h}", flush=True)
