import sys, time, numpy as np
sys.path.insert(0, '.')
import torch, stereovision_b200 as svb, oracle
from stereovision_b200.synth import make_pair
def run(H, W, D, mode, B, bs=5):
    l, r, _ = make_pair(H, W, D, seed=3)
    prm = dict(minDisparity=0, numDisparities=D, blockSize=bs, P1=8*bs*bs, P2=32*bs*bs, disp12MaxDiff=1, preFilterCap=63,
               uniquenessRatio=10, speckleWindowSize=100, speckleRange=32, mode=mode)
    ref = oracle.sgbm_compute(l, r, **prm)
    L = torch.from_numpy(l).cuda().unsqueeze(0).repeat(B, 1, 1).contiguous()
    R = torch.from_numpy(r).cuda().unsqueeze(0).repeat(B, 1, 1).contiguous()
    m = svb.StereoSGBM_create(**prm)
    t0 = time.time()
    try:
        out = m.compute(L, R); torch.cuda.synchronize()
        out = out.cpu().numpy()
        bad = sum(int((out[b] != ref).sum()) for b in range(B))
        print(f"H{H} W{W} D{D} mode{mode} B{B}: {time.time()-t0:.2f}s mismatches {bad}", flush=True)
    except Exception as e:
        print(f"H{H} W{W} D{D} mode{mode} B{B}: {time.time()-t0:.2f}s ERROR {e}", flush=True)
run(64, 200, 64, 1, 4)
run(64, 200, 64, 1, 600)
run(100, 500, 256, 1, 2)
run(100, 500, 256, 1, 100)
run(64, 200, 64, 0, 600)
