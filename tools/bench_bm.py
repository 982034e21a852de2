#!/usr/bin/env python
"""BASELINE.json configs[1]: StereoBM 640x480, numDisparities 64, SAD window 15, a batch of 256 pairs on one
B200.  Prints one JSON line: device-resident pairs/s (CUDA events), the host-buffer call (e2e), cv2.StereoBM on
the box's host cores (one pair per thread), Gdisp-evals/s.  Not the headline bench (bench.py is)."""
import argparse, ctypes, json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereovision_b200 import _lib
from stereovision_b200.synth import make_pair

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=256); ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--warmup", type=int, default=3); ap.add_argument("--H", type=int, default=480)
ap.add_argument("--W", type=int, default=640); ap.add_argument("--D", type=int, default=64)
ap.add_argument("--bs", type=int, default=15); ap.add_argument("--no-cpu", action="store_true")
a = ap.parse_args()
B, H, W, D, bs = a.batch, a.H, a.W, a.D, a.bs
pairs = [make_pair(H, W, D, seed=200 + i)[:2] for i in range(8)]
Lh = np.stack([pairs[i % 8][0] for i in range(B)]); Rh = np.stack([pairs[i % 8][1] for i in range(B)])
L = _lib.lib(); _lib.check(L.svb_init(0))
h = ctypes.c_void_p()
_lib.check(L.svb_bm_create(ctypes.byref(_lib.BmParams(0, D, bs, 31, 10, 15, 0, 0, -1, 1, 9)), ctypes.byref(h)))
Lp, Rp = torch.from_numpy(Lh).pin_memory(), torch.from_numpy(Rh).pin_memory()
Op = torch.empty((B, H, W), dtype=torch.int16).pin_memory()
Ld, Rd, Od = Lp.cuda(), Rp.cuda(), torch.empty((B, H, W), dtype=torch.int16, device="cuda")
sp = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
dev = lambda: _lib.check(L.svb_bm_compute_dev(h, Ld.data_ptr(), Rd.data_ptr(), B, H, W, Od.data_ptr(), sp))
host = lambda: _lib.check(L.svb_bm_compute_host(h, Lp.data_ptr(), Rp.data_ptr(), B, H, W, Op.data_ptr()))
for _ in range(a.warmup):
    dev()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.steps):
    dev()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.steps
host(); torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(a.steps):
    host()
e2e_ms = (time.perf_counter() - t0) / a.steps * 1e3
assert np.array_equal(Op.numpy(), Od.cpu().numpy())
cpu = None
if not a.no_cpu:
    try:
        import cv2
        from concurrent.futures import ThreadPoolExecutor
        cv2.setNumThreads(1)
        cores = len(os.sched_getaffinity(0)); n = 4 * cores
        one = lambda i: cv2.StereoBM_create(D, bs).compute(Lh[i % B], Rh[i % B])
        assert np.array_equal(one(0), Op.numpy()[0]), "GPU result differs from cv2"
        t0 = time.perf_counter()
        with ThreadPoolExecutor(cores) as ex:
            list(ex.map(one, range(n)))
        cpu = {"value": n / (time.perf_counter() - t0), "unit": "pairs/s", "cores": cores, "kind": "reference",
               "sample": f"{n} pairs, one per thread, cv2.StereoBM, cv2.setNumThreads(1)"}
    except ImportError:
        pass
ev = H * W * D / 1e9
print(json.dumps({"metric": "stereo_pairs_per_sec", "value": B / ms * 1e3, "unit": "pairs/s", "n_gpus": 1,
                  "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms, "dtype": "u16", "data": "synthetic",
                  "config": {"workload": f"bm_{W}x{H}_d{D}_bs{bs}", "pairs_per_step": B,
                             "gdisp_evals_per_s": B / ms * 1e3 * ev},
                  "e2e": {"value": B / e2e_ms * 1e3, "unit": "pairs/s", "h2d_bytes_per_step": 2 * B * H * W,
                          "d2h_bytes_per_step": 2 * B * H * W},
                  "gpu_launches": L.svb_bm_last_launches(h) * a.steps, "cpu_baseline": cpu}))
