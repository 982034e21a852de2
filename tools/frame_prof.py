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
# where a single-frame SGBM call spends its time: host-buffer call vs device call + sync vs device time (events)
import ctypes
from stereovision_b200 import _lib
L = _lib.lib()
for D, bs, mode in ((64, 5, 2), (320, 7, 2)):
    h = ctypes.c_void_p()
    _lib.check(L.svb_sgbm_create(ctypes.byref(_lib.SgbmParams(0, D, bs, 8*3*bs*bs, 32*3*bs*bs, 1, 63, 10, 100, 32, mode)), ctypes.byref(h)))
    ld, rd = torch.from_numpy(l).cuda(), torch.from_numpy(r).cuda()
    od = torch.empty((H, W), dtype=torch.int16, device="cuda")
    oh = np.empty((H, W), np.int16)
    sp = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    dev = lambda: _lib.check(L.svb_sgbm_compute_dev(h, ld.data_ptr(), rd.data_ptr(), 1, H, W, od.data_ptr(), sp))
    t(f"C compute_host D={D}", lambda: _lib.check(L.svb_sgbm_compute_host(h, l.ctypes.data, r.ctypes.data, 1, H, W, oh.ctypes.data)))
    t(f"C compute_dev+sync D={D}", lambda: (dev(), torch.cuda.synchronize()))
    t(f"C compute_dev enqueue D={D}", dev); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dev(); torch.cuda.synchronize(); e0.record()
    for _ in range(20): dev()
    e1.record(); torch.cuda.synchronize()
    print(f"device time per frame D={D}      {e0.elapsed_time(e1)/20:7.3f} ms")
    L.svb_sgbm_set_profiling(h, 1)
    for _ in range(10): dev()
    names = ctypes.create_string_buffer(32 * 16); pms = (ctypes.c_float * 16)(); pln = (ctypes.c_int * 16)()
    n = L.svb_sgbm_get_profile(h, 16, names, pms, pln)
    print({names.raw[32*i:32*i+32].split(b"\0")[0].decode(): round(pms[i] / 10, 4) for i in range(n)})
    L.svb_sgbm_set_profiling(h, 0)
