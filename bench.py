#!/usr/bin/env python
"""bench.py -- headline benchmark of the stereo hot path (BASELINE.json).

Workload (config.workload): BASELINE.json configs[3] -- StereoSGBM 1920x1080, numDisparities
256, 8-path MODE_HH, LR check + speckle filter, a batch of pairs per GPU (weak scaling:
per-GPU batch fixed as N grows).  One "step" = one pass of the SGBM hot path over one batch.

    python bench.py --gpus N --steps K --warmup W            # our arm (libsvb.so, sm_100a)
    python bench.py --impl reference --gpus N ...            # the reference's CPU path (cv2)

Prints ONE JSON line on rank 0 (see README/DESIGN.md for the keys).
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: H, W, D, mode, default batch per GPU
    "sgbm_1080p_d256_hh": dict(H=1080, W=1920, D=256, mode=1, batch=32),
    "sgbm_720p_d128_sgbm": dict(H=720, W=1280, D=128, mode=0, batch=64),
    "sgbm_480p_d64_sgbm": dict(H=480, W=640, D=64, mode=0, batch=128),
    "sgbm_480p_d320_3way": dict(H=480, W=640, D=320, mode=2, batch=64),
}


def sgbm_params(wl):
    bs = 5
    return dict(minDisparity=0, numDisparities=wl["D"], blockSize=bs, P1=8 * bs * bs, P2=32 * bs * bs,
                disp12MaxDiff=1, preFilterCap=63, uniquenessRatio=10, speckleWindowSize=100, speckleRange=32,
                mode=wl["mode"])


def b_alg_bytes(wl):
    """algorithmic bytes per pair = 2 u8 images in + s16 out + one write and one
    read of an int16 W*H*D volume = 4*W*H + 4*W*H*D."""
    return 4 * wl["W"] * wl["H"] + 4 * wl["W"] * wl["H"] * wl["D"]


def ncu_traffic_per_pair(workload):
    """DRAM bytes per pair and launch of each big kernel (dram__bytes_read.sum + dram__bytes_write.sum) from the
    committed summary of the latest `ncu --set full` capture (profiles/r02_ncu_traffic.json, written by
    tools/ncu_traffic.py from the .ncu-rep).  Returns ({kernel: bytes per pair}, source) or ({}, None)."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "r02_ncu_traffic.json")))
        if d.get("workload") != workload:
            return {}, None
        return ({k: v["dram_bytes_per_pair"] for k, v in d["kernels"].items()},
                f"profiles/r02_ncu_traffic.json ({d['source']}, {d['pairs_per_launch']} pairs per launch, scaled to this batch)")
    except Exception:
        return {}, None


def kernel_alg_bytes(wl, name):
    """Algorithmic bytes per pair of one kernel: volumes it must read / write once."""
    n2 = 2 * wl["W"] * wl["H"] * wl["D"]            # one int16 volume (upper bound: full width)
    return {"k_cost": n2, "k_sweep<first>": 2 * n2, "k_sweep<last+wta>": 2 * n2}.get(name)


def load_pinned_hashes():
    """SHA-256 of cv2's own output for the seeds this bench feeds (tests/golden/fullsize_sha256.json, written by
    tests/golden/make_golden_fullsize.py with cv2 4.13.0; only the hashes travel, never cv2)."""
    try:
        return json.load(open(os.path.join(ROOT, "tests", "golden", "fullsize_sha256.json")))["bench"]
    except Exception:
        return {}


def parity_check(workload, outs, seed0, nuniq):
    """OUTSIDE the timed region: hash the first and the last pair of every output batch in `outs`
    ({name: int16 [B,H,W] host array}) and compare with cv2's committed hash for that pair's seed."""
    import hashlib
    pinned = load_pinned_hashes()
    checked, ok = [], True
    for name, arr in outs.items():
        B = arr.shape[0]
        for i in sorted({0, B - 1}):
            seed = seed0 + i % nuniq
            want = pinned.get(f"{workload}/seed{seed}", {}).get("sha256")
            got = hashlib.sha256(np.ascontiguousarray(arr[i]).tobytes()).hexdigest()
            match = None if want is None else (got == want)
            checked.append({"output": name, "pair": i, "seed": seed, "sha256": got[:16], "match": match})
            ok = ok and bool(match)
    return {"pairs_checked": len(checked), "sha256_match": ok, "against": "tests/golden/fullsize_sha256.json "
            "(cv2 4.13.0 StereoSGBM.compute on the same seeded pairs)", "checked": checked}


def make_inputs(wl, B, seed0):
    """Synthetic pairs: a few distinct seeded scenes, tiled to the batch (generation is slow)."""
    from stereovision_b200.synth import make_pair
    nuniq = min(B, 4)
    Ls, Rs = [], []
    for i in range(nuniq):
        l, r, _ = make_pair(wl["H"], wl["W"], wl["D"], seed=seed0 + i)
        Ls.append(l); Rs.append(r)
    idx = [i % nuniq for i in range(B)]
    return np.stack([Ls[i] for i in idx]), np.stack([Rs[i] for i in idx])


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        super().__init__(daemon=True)
        self.gpu, self.samples, self.stop_flag = gpu, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [s.strip() for s in out.strip().split(",")]
                if len(f) >= 6:
                    self.samples.append(f)
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = sorted(int(s[0]) for s in self.samples if s[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(s[2 + i].lower().startswith("active") for s in self.samples)]
        mx = [int(s[1]) for s in self.samples if s[1].isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.samples)}


def bind_to_gpu_numa_node(torch, local):
    """One process per GPU: run this rank (and first-touch its pinned staging buffers) on the host cores next to
    its GPU, so that eight ranks do not pull their PCIe copies across the socket interconnect.  Returns the cpu
    list used (None when sysfs does not say)."""
    try:
        pr = torch.cuda.get_device_properties(local)
        bus = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        txt = open(f"/sys/bus/pci/devices/{bus}/local_cpulist").read().strip()
        cpus = set()
        for part in txt.split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return txt
    except Exception:
        pass
    return None


def config_of(args, wl, B):
    """The `config` object of BOTH arms (the reference arm runs on this arm's configuration; what differs -- the CPU
    sample -- is described in its cpu_baseline.sample)."""
    cfg = {"workload": args.workload, "H": wl["H"], "W": wl["W"], "numDisparities": wl["D"], "mode": wl["mode"],
           "blockSize": 5, "pairs_per_gpu_per_step": B, "l2_policy": "inputs+volumes >> L2 (126 MB)"}
    if args.scaling == "strong":
        cfg["total_pairs"] = args.total_pairs
    return cfg


def cpu_reference_run(wl, npairs, threads):
    """The reference's own CPU implementation of the path: cv2.StereoSGBM.compute
    (depth_map.py:894-909) over `npairs` distinct pairs, one pair per worker thread (cv2
    releases the GIL; MODE_SGBM/HH/3WAY are serial per image -- SURVEY 2.1).  Falls back to
    the oracle port when cv2 is not importable.  Returns (seconds, kind, cores)."""
    from concurrent.futures import ThreadPoolExecutor
    L, R = make_inputs(wl, npairs, seed0=100)
    prm = sgbm_params(wl)
    # MODE_HH keeps full C and S volumes per call: bound the worker count by free host memory
    try:
        avail = [int(l.split()[1]) * 1024 for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0]
        per_call = 6 * wl["H"] * wl["W"] * wl["D"] + (64 << 20)
        threads = max(1, min(threads, int(0.6 * avail / per_call)))
    except Exception:
        pass
    try:
        import cv2
        cv2.setNumThreads(1)
        kind = "reference"

        def one(i):
            return cv2.StereoSGBM_create(**prm).compute(L[i], R[i])
    except Exception:
        import oracle
        kind = "port"
        threads = 1

        def one(i):
            return oracle.sgbm_compute(L[i], R[i], **prm)
    t0 = time.perf_counter()
    if threads == 1:
        for i in range(npairs):
            one(i)
    else:
        with ThreadPoolExecutor(threads) as ex:
            list(ex.map(one, range(npairs)))
    return time.perf_counter() - t0, kind, threads


def cpu_single_core_ms(wl, runs=3):
    """latency of ONE pair on one host core (cv2.setNumThreads(1)), median of `runs` after the
    all-core run has warmed everything up.  None when cv2 is not importable."""
    try:
        import cv2
    except Exception:
        return None
    cv2.setNumThreads(1)
    L, R = make_inputs(wl, 1, seed0=100)
    prm = sgbm_params(wl)
    ts = []
    for _ in range(runs):
        t0 = time.perf_counter()
        cv2.StereoSGBM_create(**prm).compute(L[0], R[0])
        ts.append(1e3 * (time.perf_counter() - t0))
    return sorted(ts)[len(ts) // 2]


def run_reference_arm(args, wl, rank, world):
    if rank != 0:
        return
    cores = len(os.sched_getaffinity(0))
    sample = cores if args.ref_pairs <= 0 else args.ref_pairs
    for _ in range(args.warmup):
        cpu_reference_run(wl, min(sample, cores), cores)
    tot, kind = 0.0, "reference"
    for _ in range(args.steps):
        dt, kind, used = cpu_reference_run(wl, sample, cores)
        tot += dt
    pairs_s = sample * args.steps / tot
    unit = "pairs/s"
    line = {
        "impl": "reference", "metric": "stereo_pairs_per_sec", "value": pairs_s, "unit": unit,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot / args.steps,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "u16", "data": "synthetic",
        "config": config_of(args, wl, args.batch or wl["batch"]),
        "gdisp_evals_per_s": pairs_s * wl["H"] * wl["W"] * wl["D"] / 1e9,
        "cpu_baseline": {"value": pairs_s, "unit": unit, "cores": used, "kind": kind,
                         "sample": f"{sample} distinct pairs per step, one pair per thread, cv2.setNumThreads(1)"},
        "e2e": {"value": pairs_s, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def pageable_leg(wl, prm, Lh, Rh, steps):
    """What the reference's call site sees (depth_map.py:909): svb.StereoSGBM_create(...).compute(left, right) with
    ordinary (pageable) NumPy arrays in and a NumPy array out, a new matcher object per call like the reference's."""
    import stereovision_b200 as svb
    out = svb.StereoSGBM_create(**prm).compute(Lh, Rh)        # warm-up: workspace, staging buffers
    t0 = time.perf_counter()
    for _ in range(steps):
        out = svb.StereoSGBM_create(**prm).compute(Lh, Rh)
    return time.perf_counter() - t0, out


def fullpath_leg(wl, prm, npairs, steps, with_cpu):
    """north_star's full frame path on identical inputs: cv2.remap x 2 (BGR, CV_16SC2 maps from the synthetic
    calibration) -> cvtColor x 2 -> StereoSGBM.compute -> reprojectImageTo3D.  GPU: NumPy in / NumPy out through the
    cv2-shaped API (pageable copies inside), and tensors in / tensors out; CPU: the same calls on cv2, one pair
    per host thread."""
    import torch
    import stereovision_b200 as svb
    from stereovision_b200 import calib
    from stereovision_b200.synth import synthetic_calibration, to_bgr
    H, W = wl["H"], wl["W"]
    c = synthetic_calibration(W, H)
    R1, R2, P1, P2, Q, _, _ = calib.stereoRectify(c["mtx_left"], c["dist_left"], c["mtx_right"], c["dist_right"], (W, H),
                                                  c["R"], c["T"], alpha=0)
    ml = svb.initUndistortRectifyMap(c["mtx_left"], c["dist_left"], R1, P1, (W, H))
    mr = svb.initUndistortRectifyMap(c["mtx_right"], c["dist_right"], R2, P2, (W, H))
    Lg, Rg = make_inputs(wl, npairs, seed0=200)
    Lb, Rb = np.stack([to_bgr(x) for x in Lg]), np.stack([to_bgr(x) for x in Rg])

    def gpu(lb, rb, m_l, m_r):
        lr, rr = svb.remap(lb, m_l[0], m_l[1], svb.INTER_LINEAR), svb.remap(rb, m_r[0], m_r[1], svb.INTER_LINEAR)
        gl, gr = svb.cvtColor(lr, svb.COLOR_BGR2GRAY), svb.cvtColor(rr, svb.COLOR_BGR2GRAY)
        d = svb.StereoSGBM_create(**prm).compute(gl, gr)
        return svb.reprojectImageTo3D(d, Q)
    res = {"pairs_per_step": npairs, "stages": "remap x2 (BGR) + cvtColor x2 + StereoSGBM.compute + reprojectImageTo3D"}
    gpu(Lb, Rb, ml, mr)
    t0 = time.perf_counter()
    for _ in range(steps):
        pts = gpu(Lb, Rb, ml, mr)
    res["gpu_numpy_io_pairs_per_s"] = npairs * steps / (time.perf_counter() - t0)
    res["h2d_bytes_per_step"], res["d2h_bytes_per_step"] = int(Lb.nbytes + Rb.nbytes), int(pts.nbytes)
    dl, dr = torch.from_numpy(Lb).cuda(), torch.from_numpy(Rb).cuda()
    tml = [torch.from_numpy(m).cuda() for m in ml]
    tmr = [torch.from_numpy(m).cuda() for m in mr]
    gpu(dl, dr, tml, tmr)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        gpu(dl, dr, tml, tmr)
    torch.cuda.synchronize()
    res["gpu_device_resident_pairs_per_s"] = npairs * steps / (time.perf_counter() - t0)
    if with_cpu:
        try:
            import cv2
            from concurrent.futures import ThreadPoolExecutor
            cv2.setNumThreads(1)
            cores = len(os.sched_getaffinity(0))
            avail = [int(l.split()[1]) * 1024 for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0]
            threads = max(1, min(cores, int(0.6 * avail / (6 * H * W * wl["D"] + (64 << 20)))))

            def one(i):
                k = i % npairs
                lr = cv2.remap(Lb[k], ml[0], ml[1], cv2.INTER_LINEAR)
                rr = cv2.remap(Rb[k], mr[0], mr[1], cv2.INTER_LINEAR)
                d = cv2.StereoSGBM_create(**prm).compute(cv2.cvtColor(lr, cv2.COLOR_BGR2GRAY), cv2.cvtColor(rr, cv2.COLOR_BGR2GRAY))
                return cv2.reprojectImageTo3D(d, Q)
            t0 = time.perf_counter()
            with ThreadPoolExecutor(threads) as ex:
                ref = list(ex.map(one, range(threads)))
            res["cpu_pairs_per_s"] = threads / (time.perf_counter() - t0)
            res["cpu_cores"], res["cpu_kind"] = threads, "reference (cv2, one pair per thread)"
            # same inputs, same calls: the 3-D points must agree (disparity bit-exact -> points within 1e-4 relative)
            a, b = pts[0], ref[0]
            fin = np.isfinite(a) & np.isfinite(b)
            res["points_match_cv2"] = bool(np.array_equal(np.isfinite(a), np.isfinite(b)) and
                                           np.allclose(a[fin], b[fin], rtol=1e-4, atol=1e-4))
        except ImportError:
            res["cpu_pairs_per_s"] = None
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="svb", choices=["svb", "reference"])
    ap.add_argument("--workload", default="sgbm_1080p_d256_hh", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0, help="pairs per GPU per step (0 = workload default)")
    ap.add_argument("--ref-pairs", type=int, default=0, help="pairs per step of the CPU arm (0 = #cores)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: --batch pairs per GPU (default); strong: --total-pairs split over the GPUs (SURVEY 8(d))")
    ap.add_argument("--total-pairs", type=int, default=64, help="pairs of the whole job under --scaling strong")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the legs beside the contract's numbers (pageable NumPy e2e, full frame path, NCCL gather)")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference_arm(args, wl, rank, world)
        return

    # Libraries (NCCL's version banner) write to fd 1; the contract is ONE JSON line on stdout, so
    # everything before that line goes to stderr.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the svb arm has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = bind_to_gpu_numa_node(torch, local) if world > 1 else None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    from stereovision_b200 import _lib
    L = _lib.lib()
    _lib.check(L.svb_init(local))

    B = args.batch or wl["batch"]
    if args.scaling == "strong":                              # fixed total work: contiguous shards, first ranks take the rest
        from stereovision_b200.shard import shard_range
        a0, a1 = shard_range(args.total_pairs, rank, world)
        B = max(a1 - a0, 1)
    H, W, D = wl["H"], wl["W"], wl["D"]
    prm = sgbm_params(wl)
    Lh, Rh = make_inputs(wl, B, seed0=1000 * rank)
    # pinned host buffers for the end-to-end (host-in / host-out) leg
    Lp = torch.from_numpy(Lh).pin_memory(); Rp = torch.from_numpy(Rh).pin_memory()
    Op = torch.empty((B, H, W), dtype=torch.int16).pin_memory()
    Ld, Rd = Lp.to(dev), Rp.to(dev)
    Od = torch.empty((B, H, W), dtype=torch.int16, device=dev)

    h = ctypes.c_void_p()
    _lib.check(L.svb_sgbm_create(ctypes.byref(_lib.SgbmParams(*[prm[k] for k in (
        "minDisparity", "numDisparities", "blockSize", "P1", "P2", "disp12MaxDiff", "preFilterCap",
        "uniquenessRatio", "speckleWindowSize", "speckleRange", "mode")])), ctypes.byref(h)))
    stream = torch.cuda.current_stream()
    sp = ctypes.c_void_p(stream.cuda_stream)

    def step_dev():
        _lib.check(L.svb_sgbm_compute_dev(h, Ld.data_ptr(), Rd.data_ptr(), B, H, W, Od.data_ptr(), sp))

    def step_host():
        _lib.check(L.svb_sgbm_compute_host(h, Lp.data_ptr(), Rp.data_ptr(), B, H, W, Op.data_ptr()))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput (`value`)
    for _ in range(args.warmup):
        step_dev()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    L.svb_sgbm_set_profiling(h, 1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step_dev()
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = L.svb_sgbm_last_launches(h) * args.steps
    names = ctypes.create_string_buffer(32 * 16)
    pms = (ctypes.c_float * 16)(); pln = (ctypes.c_int * 16)()
    nst = L.svb_sgbm_get_profile(h, 16, names, pms, pln)
    L.svb_sgbm_set_profiling(h, 0)
    stages = {names.raw[32 * i:32 * i + 32].split(b"\0")[0].decode(): (pms[i], pln[i]) for i in range(max(nst, 0))}

    # ---- end-to-end through the host-buffer C-ABI call (`e2e`): H2D + compute + D2H each step
    step_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_host()
    barrier()
    e2e_s = time.perf_counter() - t0
    sampler.stop_flag = True
    sampler.join(timeout=2)

    # ---- parity of what was just timed (outside the timed region): first and last pair of the device-resident
    # output and of the host-path output against cv2's committed hashes; the wavefront flag of the device path
    _lib.check(L.svb_sgbm_check(h))
    parity = parity_check(args.workload, {"value(dev)": Od.cpu().numpy(), "e2e(host)": Op.numpy()},
                          seed0=1000 * rank, nuniq=min(B, 4))
    # ---- legs beside the contract's numbers (all outside the timed regions above)
    extras = {}
    if world > 1 and not args.no_extras:
        # the one collective this path has: gather every rank's disparities (NCCL), check what arrived
        from stereovision_b200.shard import gather_outputs, shard_sizes
        import hashlib
        sizes = [B] * world if args.scaling == "weak" else shard_sizes(args.total_pairs, world)
        total = sum(sizes)
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        gather_outputs(Od, total)                               # warm-up (communicator set-up)
        barrier()
        g0.record(stream)
        allout = gather_outputs(Od, total)
        g1.record(stream)
        barrier()
        extras["gather_gbs"] = allout.numel() * 2 * (world - 1) / world / (g0.elapsed_time(g1) / 1e3) / 1e9
        pinned = load_pinned_hashes()
        ok, off = True, 0
        for r, n in enumerate(sizes):                           # pair 0 of every rank's shard, as it arrived HERE
            want = pinned.get(f"{args.workload}/seed{1000 * r}", {}).get("sha256")
            got = hashlib.sha256(allout[off].cpu().numpy().tobytes()).hexdigest()
            ok = ok and want is not None and got == want
            off += n
        extras["gathered_match"] = ok
        del allout
    L.svb_sgbm_destroy(h)
    h = None
    if rank == 0 and world == 1 and not args.no_extras:
        dt, outp = pageable_leg(wl, prm, Lh, Rh, min(args.steps, 3))
        extras["e2e_pageable"] = {"value": B * min(args.steps, 3) / dt, "unit": "pairs/s",
                                  "call": "svb.StereoSGBM_create(**prm).compute(np_left, np_right), pageable NumPy in / NumPy out",
                                  "h2d_bytes_per_step": 2 * B * H * W, "d2h_bytes_per_step": 2 * B * H * W,
                                  "equals_pinned_path": bool(np.array_equal(outp, Op.numpy()))}
        import stereovision_b200 as svb_pkg
        svb_pkg.release_workspaces()
        extras["fullpath"] = fullpath_leg(wl, prm, 8, 2, not args.no_cpu_baseline)
        svb_pkg.release_workspaces()
    tmax = torch.tensor([ms, e2e_s * 1e3, 0.0 if parity["sha256_match"] and extras.get("gathered_match", True) else 1.0],
                        dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms, e2e_ms = float(tmax[0]), float(tmax[1])
    parity["all_ranks_match"] = float(tmax[2]) == 0.0

    if rank == 0:
        pairs = (B * world if args.scaling == "weak" else args.total_pairs) * args.steps
        value = pairs / (ms / 1e3)
        e2e = pairs / (e2e_ms / 1e3)
        balg = b_alg_bytes(wl)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        # dominant stage: all launches of the kernel with the largest share of the step
        dom = max(stages, key=lambda k: stages[k][0]) if stages else None
        sgbm_ms = sum(v[0] for v in stages.values())
        roof = None
        if dom:
            dms, dl = stages[dom]
            # the hot path as one unit: B_alg x pairs over the summed device time of its kernels
            ach = balg * B * args.steps / (sgbm_ms / 1e3) / 1e9
            traffic, traffic_src = ncu_traffic_per_pair(args.workload)
            tr = traffic.get(dom)
            kalg = kernel_alg_bytes(wl, dom)
            roof = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                    "traffic": tr * B if tr else None,
                    "traffic_source": ("ncu dram bytes of the dominant kernel per launch: " + traffic_src) if tr else None,
                    "dominant_kernel_alg_bytes_per_launch": kalg * B if kalg else None,
                    "dominant_kernel_achieved": (kalg * B / (dms / max(dl, 1) / 1e3) / 1e9) if kalg else None,
                    "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650",
                    "kernel": "sgbm chain (all stages)", "dominant_kernel": dom,
                    "dominant_share": dms / sgbm_ms if sgbm_ms else None,
                    "dominant_ms_per_launch": dms / max(dl, 1),
                    "b_alg_bytes_per_pair": balg,
                    "stages_ms_per_step": {k: v[0] / args.steps for k, v in stages.items()}}
        line = {
            "metric": "stereo_pairs_per_sec", "value": value, "unit": "pairs/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "u16",
            "data": "synthetic",
            "config": config_of(args, wl, B), "host_numa_cpus": numa,
            "gdisp_evals_per_s": value * H * W * D / 1e9,
            "e2e": {"value": e2e, "unit": "pairs/s", "h2d_bytes_per_step": 2 * H * W * (pairs // args.steps),
                    "d2h_bytes_per_step": 2 * H * W * (pairs // args.steps)},
            "gpu_launches": launches * world,
            "parity": parity,
            "multi_gpu_identical": (parity["all_ranks_match"] if world > 1 else None),
            "clocks": sampler.summary(),
            "roofline": roof,
        }
        line.update(extras)
        if args.scaling == "strong":
            line["strong_scaling_note"] = ("fixed total work: fewer pairs per launch leave the wavefront's ramp-up and "
                                                     "ramp-down (one band per SM at a time) a larger share of each sweep")
        if world == 1 and not args.no_cpu_baseline:
            cores = len(os.sched_getaffinity(0))
            dt, kind, used = cpu_reference_run(wl, cores, cores)
            line["cpu_baseline"] = {"value": cores / dt, "unit": "pairs/s", "cores": used, "kind": kind,
                                    "sample": f"{cores} distinct pairs once, one pair per thread (cv2.setNumThreads(1))",
                                    "single_core_ms_per_pair": cpu_single_core_ms(wl)}
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        print(json.dumps(line), flush=True)
    if h is not None:
        L.svb_sgbm_destroy(h)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
