#!/usr/bin/env python
"""Device-path time of one StereoSGBM configuration with the per-stage profile:  python tools/time_cfg.py H W D mode bs B
(with SVB_LIBRARY=build/libsvb_<tag>.so to A/B a kernel variant)."""
import ctypes, hashlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import stereovision_b200 as svb
from stereovision_b200 import _lib
from stereovision_b200.synth import make_pair

H, W, D, mode, bs, B = (int(v) for v in sys.argv[1:7])
pairs = [make_pair(H, W, min(D, 256), seed=s) for s in range(4)]
L = torch.from_numpy(np.stack([pairs[i % 4][0] for i in range(B)])).cuda()
R = torch.from_numpy(np.stack([pairs[i % 4][1] for i in range(B)])).cuda()
m = svb.StereoSGBM_create(minDisparity=0, numDisparities=D, blockSize=bs, P1=8 * bs * bs, P2=32 * bs * bs, disp12MaxDiff=1,
                          preFilterCap=63, uniquenessRatio=10, speckleWindowSize=100, speckleRange=32, mode=mode)
out = m.compute(L, R)
torch.cuda.synchronize()
lib = _lib.lib()
lib.svb_sgbm_set_profiling(m._last, 1)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    out = m.compute(L, R)
e1.record()
torch.cuda.synchronize()
names = ctypes.create_string_buffer(16 * 32)
ms = (ctypes.c_float * 16)()
launches = (ctypes.c_int * 16)()
n = lib.svb_sgbm_get_profile(m._last, 16, names, ms, launches)
prof = {names.raw[32 * i:32 * i + 32].split(b"\0")[0].decode(): round(ms[i] / 5, 2) for i in range(max(n, 0))}
t = e0.elapsed_time(e1) / 5
print(f"{H}x{W} D{D} mode{mode} bs{bs} B{B}: {t:8.3f} ms {B / t * 1e3:9.1f} pairs/s sha {hashlib.sha256(out.cpu().numpy().tobytes()).hexdigest()[:10]} {prof}")
