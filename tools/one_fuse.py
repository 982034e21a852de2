#!/usr/bin/env python
"""ncu driver for BASELINE config 5: a few fuse_depth_maps calls at 1080p (k_fill_rows, k_blend, k_bilat_lut, k_bilateral)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import stereovision_b200 as svb
from stereovision_b200.synth import make_pair, make_aux_depths

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
_, _, gt = make_pair(1080, 1920, 256, seed=1)
midas, flow = make_aux_depths(gt, 1)
rep = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda().unsqueeze(0).repeat(B, 1, 1).contiguous()
dispf = rep(gt.astype(np.float32))
norm, conf = svb.normalize_disparity(dispf, 0, 256)
cal = svb.calibrate_midas_to_stereo(rep(midas), dispf, conf)
fl = svb.normalize_to_stereo_range(rep(flow), dispf)
for _ in range(3):
    out = svb.fuse_depth_maps(norm, conf, cal, None, fl, True)
torch.cuda.synchronize()
print("ok", tuple(out.shape), int(out.sum()))
