"""Per-row device timing of the stages around the SGBM matcher: remap(+gray),
s16 -> f32 + median5 + depth, reprojectImageTo3D, the fusion chain (a13-a16), occlusion statistics and
StereoBM.  CUDA events on the current stream, inputs resident in HBM and larger than L2 in total.
Prints one JSON object per row:  python tools/bench_stages.py [--H 1080 --W 1920 --batch 16]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--H", type=int, default=1080)
    ap.add_argument("--W", type=int, default=1920)
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--iters", type=int, default=5)
    a = ap.parse_args()
    import torch
    import stereovision_b200 as svb
    from stereovision_b200.synth import make_pair, make_aux_depths, to_bgr
    H, W, B = a.H, a.W, a.batch
    peak = 6551.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    dev = torch.device("cuda", 0)
    l, r, gt = make_pair(H, W, 256, seed=1)
    midas, flow = make_aux_depths(gt, 1)
    rep = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev).unsqueeze(0).repeat(B, *([1] * x.ndim)).contiguous()
    bgr = rep(to_bgr(l))
    gray_l, gray_r = rep(l), rep(r)
    # rectification maps from the synthetic calibration through the package's own path (host stereoRectify + device maps)
    from stereovision_b200 import calib
    from stereovision_b200.synth import synthetic_calibration
    sc = synthetic_calibration(W, H)
    R1, _, P1, _, Q, _, _ = calib.stereoRectify(sc["mtx_left"], sc["dist_left"], sc["mtx_right"], sc["dist_right"], (W, H),
                                                sc["R"], sc["T"], alpha=0)
    m1, m2 = svb.initUndistortRectifyMap(sc["mtx_left"], sc["dist_left"], R1, P1, (W, H), as_tensor=True)
    disp16 = rep((gt * 16).astype(np.int16))
    dispf = disp16.float() / 16.0
    midas_d, flow_d = rep(midas), rep(flow)
    norm, conf = svb.normalize_disparity(dispf, 0, 256)
    cal_d = svb.calibrate_midas_to_stereo(midas_d, dispf, conf)
    flow_n = svb.normalize_to_stereo_range(flow_d, dispf)
    bm = svb.StereoBM_create(64, 15)
    px = H * W

    def chain():
        n_, c_ = svb.normalize_disparity(dispf, 0, 256)
        m_ = svb.calibrate_midas_to_stereo(midas_d, dispf, c_)
        f_ = svb.normalize_to_stereo_range(flow_d, dispf)
        return svb.fuse_depth_maps(n_, c_, m_, None, f_, True)

    Kc = np.array([[0.78 * W, 0, W / 2 + 3], [0, 0.78 * W, H / 2 - 2], [0, 0, 1]], np.float64)
    dc = np.array([-0.10, 0.02, 1e-3, -1e-3, 0], np.float64)
    gen = torch.Generator(device=dev).manual_seed(1)
    pred_d = torch.randn((B, 384, 384), device=dev, generator=gen) * 200 + 500
    flow2_d = torch.randn((B, H, W, 2), device=dev, generator=gen) * 3
    stab_d = torch.rand((B, H, W), device=dev, generator=gen) + 0.1
    Hm = np.array([[1.01, 0.002, 1.5], [-0.003, 0.99, -2.0], [1e-6, -2e-6, 1.0]])
    rows = [
        ("initUndistortRectifyMap (N1, per map pair)", lambda: svb.initUndistortRectifyMap(Kc, dc, None, Kc, (W, H), as_tensor=True), 6 * px / B),
        ("resize BGR x0.33 (a1, fused_depth_map.py:473-476)", lambda: svb.resize(bgr, (int(W * 0.33), int(H * 0.33))), 3 * px + 3 * int(W * 0.33) * int(H * 0.33)),
        ("cvtColor BGR2GRAY (a2)", lambda: svb.cvtColor(bgr), 4 * px),
        ("remap BGR + gray (a1,a2)", lambda: svb.remap(bgr, m1, m2, svb.INTER_LINEAR, to_gray=True), (3 + 1) * px),
        ("remap BGR (a1)", lambda: svb.remap(bgr, m1, m2, svb.INTER_LINEAR), 6 * px),
        ("s16->f32 + median5 + depth (a10,a11)", lambda: svb.disparity_to_depth(disp16), 10 * px),
        ("reprojectImageTo3D s16 (a12)", lambda: svb.reprojectImageTo3D(disp16, Q), 14 * px),
        ("normalise + confidence (a13)", lambda: svb.normalize_disparity(dispf, 0, 256), 12 * px),
        ("calibrate_midas_to_stereo (a14)", lambda: svb.calibrate_midas_to_stereo(midas_d, dispf, conf), 16 * px),
        ("normalize_to_stereo_range (a15)", lambda: svb.normalize_to_stereo_range(flow_d, dispf), 12 * px),
        ("fuse_depth_maps (a16)", lambda: svb.fuse_depth_maps(norm, conf, cal_d, None, flow_n, True), 17 * px),
        ("fusion chain a13-a16 (C5)", chain, 17 * px),
        ("occlusion statistics (a17)", lambda: svb.occlusion_stats(gray_l), px),
        ("MiDaS upsample 384x384 -> frame + normalise (N3)", lambda: svb.upsample_normalize(pred_d, (H, W)), 12 * px),
        ("contrast confidence u8 (N3)", lambda: svb.contrast_confidence(gray_l), 4 * px),
        ("flow residual depth + bilateral (N3)", lambda: svb.flow_residual_depth(flow2_d, Hm, stab_d, True), 36 * px),
        ("bilateralFilter f32 9/75/75 (N3)", lambda: svb.bilateralFilter(flow_d, 9, 75, 75), 12 * px),
        ("StereoBM 64/15 (a18)", lambda: bm.compute(gray_l, gray_r), 4 * px),
    ]
    for name, fn, bytes_per_item in rows:
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.iters
        gbs = bytes_per_item * B / (ms / 1e3) / 1e9
        print(json.dumps({"row": name, "H": H, "W": W, "batch": B, "ms_per_batch": round(ms, 4),
                          "items_per_s": round(B / (ms / 1e3), 1), "alg_bytes_per_item": bytes_per_item,
                          "achieved_GBps": round(gbs, 1), "hbm_peak_GBps": peak, "frac": round(gbs / peak, 4)}), flush=True)


if __name__ == "__main__":
    main()
