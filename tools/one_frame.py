import sys, numpy as np
sys.path.insert(0, "/root/repo")
import stereovision_b200 as svb
from stereovision_b200.synth import make_pair
l, r, _ = make_pair(480, 640, 64, seed=9)
st = svb.StereoSGBM_create(minDisparity=0, numDisparities=64, blockSize=5, P1=600, P2=2400, disp12MaxDiff=1, uniquenessRatio=10, speckleWindowSize=100, speckleRange=32, preFilterCap=63, mode=2)
for _ in range(3): st.compute(l, r)
