"""Where the host-buffer call's time goes (1080p / D256 / HH, 32 pairs): copies, kernels, fixed host work."""
import ctypes, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereovision_b200 import _lib
from bench import WORKLOADS, sgbm_params, make_inputs

wl = WORKLOADS["sgbm_1080p_d256_hh"]; B, H, W = int(os.environ.get("B", 32)), wl["H"], wl["W"]
L = _lib.lib(); _lib.check(L.svb_init(0))
Lh, Rh = make_inputs(wl, B, 0)
Lp, Rp = torch.from_numpy(Lh).pin_memory(), torch.from_numpy(Rh).pin_memory()
Op = torch.empty((B, H, W), dtype=torch.int16).pin_memory()
Ld, Rd, Od = Lp.cuda(), Rp.cuda(), torch.empty((B, H, W), dtype=torch.int16, device="cuda")
h = ctypes.c_void_p()
prm = sgbm_params(wl)
_lib.check(L.svb_sgbm_create(ctypes.byref(_lib.SgbmParams(*[prm[k] for k in (
    "minDisparity", "numDisparities", "blockSize", "P1", "P2", "disp12MaxDiff", "preFilterCap",
    "uniquenessRatio", "speckleWindowSize", "speckleRange", "mode")])), ctypes.byref(h)))
sp = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

def t(fn, n=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn(); torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3

print("H2D 2 x %.0f MB      %.2f ms" % (Lh.nbytes / 1e6, t(lambda: (Ld.copy_(Lp, non_blocking=True), Rd.copy_(Rp, non_blocking=True)))))
print("D2H %.0f MB          %.2f ms" % (Op.numel() * 2 / 1e6, t(lambda: Op.copy_(Od, non_blocking=True))))
print("compute_dev + sync  %.2f ms" % t(lambda: _lib.check(L.svb_sgbm_compute_dev(h, Ld.data_ptr(), Rd.data_ptr(), B, H, W, Od.data_ptr(), sp))))
t0 = time.perf_counter()
for _ in range(5):
    _lib.check(L.svb_sgbm_compute_dev(h, Ld.data_ptr(), Rd.data_ptr(), B, H, W, Od.data_ptr(), sp))
host_ms = (time.perf_counter() - t0) / 5 * 1e3
torch.cuda.synchronize()
print("compute_dev host-side enqueue time %.2f ms" % host_ms)
print("compute_host        %.2f ms" % t(lambda: _lib.check(L.svb_sgbm_compute_host(h, Lp.data_ptr(), Rp.data_ptr(), B, H, W, Op.data_ptr()))))
