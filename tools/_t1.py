import numpy as np, sys
sys.path.insert(0,'.')
import stereovision_b200 as svb, oracle
from stereovision_b200.synth import make_pair
l,r,_=make_pair(60,96,16,seed=11)
prm=dict(minDisparity=0,numDisparities=16,blockSize=3,P1=72,P2=288,disp12MaxDiff=1,preFilterCap=63,uniquenessRatio=10,speckleWindowSize=100,speckleRange=32,mode=0)
got=svb.StereoSGBM_create(**prm).compute(l,r)
ref=oracle.sgbm_compute(l,r,**prm)
print("mismatch",(got!=ref).sum())
