#!/usr/bin/env python
"""ncu driver: a few device-resident steps of a bench.py workload and nothing else.
    python tools/one_step.py [workload] [pairs] [steps]
e.g. ncu --set full -k regex:'k_sweep|k_cost' -s 3 -c 3 ... python tools/one_step.py sgbm_1080p_d256_hh 32 2"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402
from stereovision_b200 import _lib  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "sgbm_1080p_d256_hh"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
wl = bench.WORKLOADS[name]
prm = bench.sgbm_params(wl)
L = _lib.lib()
_lib.check(L.svb_init(0))
Lh, Rh = bench.make_inputs(wl, B, 0)
Ld, Rd = torch.from_numpy(Lh).cuda(), torch.from_numpy(Rh).cuda()
Od = torch.empty((B, wl["H"], wl["W"]), dtype=torch.int16, device="cuda")
h = ctypes.c_void_p()
order = ("minDisparity", "numDisparities", "blockSize", "P1", "P2", "disp12MaxDiff", "preFilterCap", "uniquenessRatio",
         "speckleWindowSize", "speckleRange", "mode")
_lib.check(L.svb_sgbm_create(ctypes.byref(_lib.SgbmParams(*[prm[k] for k in order])), ctypes.byref(h)))
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
for _ in range(steps):
    _lib.check(L.svb_sgbm_compute_dev(h, Ld.data_ptr(), Rd.data_ptr(), B, wl["H"], wl["W"], Od.data_ptr(), st))
torch.cuda.synchronize()
_lib.check(L.svb_sgbm_check(h))
print("ok", name, B, steps, int((Od >= 0).sum()))
L.svb_sgbm_destroy(h)
