"""GPU: the boundary contract around the kernels: calls from worker threads, successive computes of
one handle on different streams, the wavefront's error flag on the device path, the handle pool's eviction, and
the caller's current device."""
import ctypes
import threading
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

import oracle
import stereovision_b200 as svb
from stereovision_b200 import _lib
from stereovision_b200.synth import make_pair

pytestmark = pytest.mark.gpu

PRM = dict(minDisparity=0, numDisparities=64, blockSize=5, P1=200, P2=800, disp12MaxDiff=1, preFilterCap=63,
           uniquenessRatio=10, speckleWindowSize=100, speckleRange=32, mode=1)
ORDER = ("minDisparity", "numDisparities", "blockSize", "P1", "P2", "disp12MaxDiff", "preFilterCap",
         "uniquenessRatio", "speckleWindowSize", "speckleRange", "mode")


def _torch():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch


def test_compute_from_pool_worker_while_main_thread_computes():
    """fused_depth_map.py:2591-2598: create_depth_map_stereo_scaled runs in a ThreadPoolExecutor worker while the
    main thread does other work on the same GPU."""
    _torch()
    H, W = 120, 300
    pairs = [make_pair(H, W, 64, seed=200 + i)[:2] for i in range(4)]
    refs = [oracle.sgbm_compute(l, r, **PRM) for l, r in pairs]
    stop = threading.Event()

    def worker(i):
        out = []
        for _ in range(6):
            l, r = pairs[i]
            out.append(svb.StereoSGBM_create(**PRM).compute(l, r))
        return out

    l0, r0 = pairs[0]
    with ThreadPoolExecutor(2) as ex:
        futs = [ex.submit(worker, 1), ex.submit(worker, 2)]
        mine = []
        while not all(f.done() for f in futs):
            mine.append(svb.StereoSGBM_create(**PRM).compute(l0, r0))
            svb.medianBlur(mine[-1], 3)
        stop.set()
        for i, f in zip((1, 2), futs):
            for got in f.result():
                assert np.array_equal(got, refs[i])
    assert mine and all(np.array_equal(g, refs[0]) for g in mine)


def test_same_handle_on_two_streams_is_ordered():
    """Two computes with the same parameters share one handle and one workspace.  Enqueued on different torch
    streams they must not overlap on the device (the handle chains them with an event)."""
    torch = _torch()
    H, W, B = 160, 400, 24
    la, ra, _ = make_pair(H, W, 64, seed=301)
    lb, rb, _ = make_pair(H, W, 64, seed=302)
    ref_a, ref_b = oracle.sgbm_compute(la, ra, **PRM), oracle.sgbm_compute(lb, rb, **PRM)
    A = [torch.from_numpy(x).cuda().unsqueeze(0).repeat(B, 1, 1).contiguous() for x in (la, ra)]
    Bt = [torch.from_numpy(x).cuda().unsqueeze(0).repeat(B, 1, 1).contiguous() for x in (lb, rb)]
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    torch.cuda.synchronize()
    m = svb.StereoSGBM_create(**PRM)
    outs = []
    for it in range(6):
        with torch.cuda.stream(s1):
            oa = m.compute(*A)
        with torch.cuda.stream(s2):
            ob = m.compute(*Bt)
        outs.append((oa, ob))
    torch.cuda.synchronize()
    m.check()
    ta, tb = torch.from_numpy(ref_a).cuda(), torch.from_numpy(ref_b).cuda()
    for oa, ob in outs:
        assert int((oa != ta.unsqueeze(0)).sum()) == 0 and int((ob != tb.unsqueeze(0)).sum()) == 0


def test_device_path_reports_the_sweep_error_flag():
    """svb_sgbm_compute_dev does not synchronise, so a tripped wavefront flag surfaces (a) through svb_sgbm_check
    and (b) at the next compute on the handle once the flagged one has finished.  The flag is raised by hand."""
    torch = _torch()
    lib = _lib.lib()
    H, W = 64, 200
    l, r, _ = make_pair(H, W, 64, seed=5)
    L, R = torch.from_numpy(l).cuda(), torch.from_numpy(r).cuda()
    out = torch.empty((H, W), dtype=torch.int16, device="cuda")
    h = ctypes.c_void_p()
    _lib.check(lib.svb_sgbm_create(ctypes.byref(_lib.SgbmParams(*[PRM[k] for k in ORDER])), ctypes.byref(h)))
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def run():
        return lib.svb_sgbm_compute_dev(h, L.data_ptr(), R.data_ptr(), 1, H, W, out.data_ptr(), st)
    try:
        assert run() == 0 and lib.svb_sgbm_check(h) == 0
        assert np.array_equal(out.cpu().numpy(), oracle.sgbm_compute(l, r, **PRM))
        # (a) explicit check
        assert lib.svb_sgbm_debug_trip_error(h) == 0
        assert run() == 0
        assert lib.svb_sgbm_check(h) == -3 and b"spin limit" in lib.svb_last_error()
        assert lib.svb_sgbm_check(h) == 0                     # reported once, cleared
        assert run() == 0 and lib.svb_sgbm_check(h) == 0
        # (b) the next compute reports a finished predecessor's flag without anybody calling check
        assert lib.svb_sgbm_debug_trip_error(h) == 0
        assert run() == 0
        torch.cuda.synchronize()
        assert run() == -3
        assert run() == 0 and lib.svb_sgbm_check(h) == 0
        assert np.array_equal(out.cpu().numpy(), oracle.sgbm_compute(l, r, **PRM))
    finally:
        lib.svb_sgbm_destroy(h)
    # the cv2-shaped surface: check() raises
    m = svb.StereoSGBM_create(**PRM)
    m.compute(L, R)
    _lib.check(lib.svb_sgbm_debug_trip_error(m._last))
    m.compute(L, R)
    with pytest.raises(_lib.SvbError):
        m.check()
    m.check()


def test_handle_pool_evicts_least_recently_used():
    """The reference changes matcher parameters live (fused_depth_map.py:3241-3289): every new tuple gets a handle,
    the pool keeps a bounded number per thread and frees the rest."""
    torch = _torch()
    svb.release_workspaces()
    l, r, _ = make_pair(48, 160, 32, seed=9)
    free0 = torch.cuda.mem_get_info()[0]
    for k in range(10):
        prm = dict(PRM, numDisparities=32, uniquenessRatio=k, mode=0)
        got = svb.StereoSGBM_create(**prm).compute(l, r)
        if k in (0, 9):
            assert np.array_equal(got, oracle.sgbm_compute(l, r, **prm))
    from stereovision_b200 import cv_api
    assert svb.pool_size() <= cv_api._POOL_PER_THREAD
    # a finished thread's handles go away with the next insertion from any thread
    t = threading.Thread(target=lambda: svb.StereoSGBM_create(**dict(PRM, uniquenessRatio=33)).compute(l, r))
    t.start(); t.join()
    svb.StereoSGBM_create(**dict(PRM, uniquenessRatio=34)).compute(l, r)
    with cv_api._pool_lock:
        assert all(k[0] == threading.get_ident() for k in cv_api._pool)
    svb.release_workspaces()
    assert svb.pool_size() == 0
    assert torch.cuda.mem_get_info()[0] >= free0 - (64 << 20)


def test_callers_current_device_is_restored():
    torch = _torch()
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    l, r, _ = make_pair(48, 160, 32, seed=9)
    prm = dict(PRM, numDisparities=32)
    ref = oracle.sgbm_compute(l, r, **prm)
    torch.cuda.set_device(0)
    L1, R1 = torch.from_numpy(l).to("cuda:1"), torch.from_numpy(r).to("cuda:1")
    out = svb.StereoSGBM_create(**prm).compute(L1, R1)
    assert torch.cuda.current_device() == 0 and out.device.index == 1
    with torch.cuda.device(1):
        out1 = svb.StereoSGBM_create(**prm).compute(L1, R1)
    out0 = svb.StereoSGBM_create(**prm).compute(L1.to("cuda:0"), R1.to("cuda:0"))
    out1b = svb.StereoSGBM_create(**prm).compute(L1, R1)      # cuda:1 data again while device 0 is current
    assert torch.cuda.current_device() == 0
    for o in (out, out1, out0, out1b):
        assert np.array_equal(o.cpu().numpy(), ref)


def test_pitched_inputs_without_a_copy():
    """SURVEY 8(b) row_pitch / img_pitch: a crop of a larger frame (rows not adjacent in memory, images not adjacent in
    the batch) goes through svb_sgbm_compute_pitched_dev as it is and gives the bytes of the dense copy; gray and BGR,
    the raw C ABI and the cv2-shaped surface (which used to .contiguous() every view)."""
    torch = _torch()
    lib = _lib.lib()
    H, W, B = 60, 200, 3
    big_l = torch.randint(0, 256, (B, H + 13, W + 24), dtype=torch.uint8, device="cuda")
    big_r = torch.randint(0, 256, (B, H + 13, W + 24), dtype=torch.uint8, device="cuda")
    for b in range(B):
        l, r, _ = make_pair(H, W, 64, seed=400 + b)
        big_l[b, 5:5 + H, 8:8 + W] = torch.from_numpy(l).cuda()
        big_r[b, 5:5 + H, 8:8 + W] = torch.from_numpy(r).cuda()
    vl, vr = big_l[:, 5:5 + H, 8:8 + W], big_r[:, 5:5 + H, 8:8 + W]
    assert not vl.is_contiguous()
    m = svb.StereoSGBM_create(**PRM)
    want = m.compute(vl.contiguous(), vr.contiguous())
    got = m.compute(vl, vr)                                    # the pitched path
    assert torch.equal(got, want)
    assert np.array_equal(got[1].cpu().numpy(), oracle.sgbm_compute(vl[1].cpu().numpy(), vr[1].cpu().numpy(), **PRM))
    # one image of the batch, every second image (img_pitch = 2 frames)
    assert torch.equal(m.compute(vl[2], vr[2]), want[2])
    assert torch.equal(m.compute(vl[::2], vr[::2]), want[::2])
    # raw C ABI, pitches in bytes; too small a pitch is rejected
    h = ctypes.c_void_p()
    _lib.check(lib.svb_sgbm_create(ctypes.byref(_lib.SgbmParams(*[PRM[k] for k in ORDER])), ctypes.byref(h)))
    out = torch.empty((B, H, W), dtype=torch.int16, device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.svb_sgbm_compute_pitched_dev(h, vl.data_ptr(), vr.data_ptr(), B, H, W, 1, vl.stride(1), vl.stride(0), out.data_ptr(), st)
    assert rc == 0 and lib.svb_sgbm_check(h) == 0 and torch.equal(out, want)
    assert lib.svb_sgbm_compute_pitched_dev(h, vl.data_ptr(), vr.data_ptr(), B, H, W, 1, W - 1, 0, out.data_ptr(), st) == -1
    lib.svb_sgbm_destroy(h)
    # BGR crop
    cl = torch.randint(0, 256, (H + 4, W + 6, 3), dtype=torch.uint8, device="cuda")
    cr = torch.roll(cl, -5, dims=1)
    wl, wr = cl[2:2 + H, 3:3 + W], cr[2:2 + H, 3:3 + W]
    prm3 = dict(PRM, P1=600, P2=2400)
    assert torch.equal(svb.StereoSGBM_create(**prm3).compute(wl, wr), svb.StereoSGBM_create(**prm3).compute(wl.contiguous(), wr.contiguous()))


def test_pageable_numpy_batches_go_through_pinned_staging():
    """The reference's call site hands NumPy (pageable) arrays to compute() (depth_map.py:909).  Batches of them are
    copied to pinned staging by the library's host threads while the first sub-ranges already run, and copied out per
    downloaded sub-range (svb_sgbm_compute_cn_host); the bytes must be those of the device path -- also when the
    workspace limit forces several chunks, for colour input, and for arrays that were pinned by the caller."""
    torch = _torch()
    H, W, B = 480, 640, 9                                     # 9 x (2 x 0.3 + 0.6) MB > the 8 MB threshold; odd batch
    L = np.stack([make_pair(H, W, 64, seed=700 + b)[0] for b in range(B)])
    R = np.stack([make_pair(H, W, 64, seed=700 + b)[1] for b in range(B)])
    m = svb.StereoSGBM_create(**PRM)
    want = m.compute(torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()).cpu().numpy()
    got = m.compute(L, R)
    assert got.dtype == np.int16 and np.array_equal(got, want)
    assert np.array_equal(got[4], oracle.sgbm_compute(L[4], R[4], **PRM))
    # several chunks: a workspace limit of ~3 pairs
    lib = _lib.lib()
    h = ctypes.c_void_p()
    _lib.check(lib.svb_sgbm_create(ctypes.byref(_lib.SgbmParams(*[PRM[k] for k in ORDER])), ctypes.byref(h)))
    per = lib.svb_sgbm_workspace_bytes(h, 1, H, W)
    _lib.check(lib.svb_sgbm_set_workspace_limit(h, int(3.5 * per)))
    out = np.empty((B, H, W), np.int16)
    _lib.check(lib.svb_sgbm_compute_cn_host(h, L.ctypes.data, R.ctypes.data, B, H, W, 1, out.ctypes.data))
    assert np.array_equal(out, want)
    # pinned caller buffers take the direct path; same bytes
    pl, pr = torch.from_numpy(L).pin_memory(), torch.from_numpy(R).pin_memory()
    po = torch.empty((B, H, W), dtype=torch.int16).pin_memory()
    _lib.check(lib.svb_sgbm_compute_cn_host(h, pl.data_ptr(), pr.data_ptr(), B, H, W, 1, po.data_ptr()))
    assert np.array_equal(po.numpy(), want)
    lib.svb_sgbm_destroy(h)
    # colour (BGR) batch
    Lc, Rc = np.repeat(L[:5, :, :, None], 3, axis=3), np.repeat(R[:5, :, :, None], 3, axis=3)
    prm3 = dict(PRM, P1=600, P2=2400)
    m3 = svb.StereoSGBM_create(**prm3)
    assert np.array_equal(m3.compute(Lc, Rc), m3.compute(torch.from_numpy(Lc).cuda(), torch.from_numpy(Rc).cuda()).cpu().numpy())


def test_pageable_numpy_batches_stereobm():
    """The same staging for svb_bm_compute_host: a NumPy batch above the threshold gives the bytes of the device path."""
    torch = _torch()
    H, W, B = 480, 640, 11
    L = np.stack([make_pair(H, W, 64, seed=800 + b)[0] for b in range(B)])
    R = np.stack([make_pair(H, W, 64, seed=800 + b)[1] for b in range(B)])
    bm = svb.StereoBM_create(numDisparities=64, blockSize=15)
    want = bm.compute(torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()).cpu().numpy()
    assert np.array_equal(bm.compute(L, R), want)
    assert np.array_equal(want[3], oracle.bm_compute(L[3], R[3], numDisparities=64, blockSize=15))


def test_staged_copies_round_trip_numpy_arrays():
    """svb_copy_h2d / svb_copy_d2h (pinned staging + host threads for pageable NumPy arrays): sizes around the chunk
    size, several dtypes, a pinned source, and the public calls that use them (a 1080p BGR cvtColor)."""
    torch = _torch()
    from stereovision_b200 import cv_api
    rng = np.random.default_rng(5)
    for n in ((4 << 20) + 1, (8 << 20) - 3, (8 << 20), (8 << 20) + 4099, (33 << 20) + 7):
        a = rng.integers(0, 256, n, dtype=np.uint8)
        t = cv_api._upload(a)
        assert t.is_cuda and t.dtype == torch.uint8 and np.array_equal(t.cpu().numpy(), a)
        assert np.array_equal(cv_api._download(t), a)
    for dt in (np.int16, np.float32, np.float64):
        a = rng.standard_normal((3, 700, 1100)).astype(dt) if dt != np.int16 else rng.integers(-3000, 3000, (3, 700, 1100)).astype(dt)
        b = cv_api._download(cv_api._upload(a))
        assert b.dtype == a.dtype and b.shape == a.shape and np.array_equal(b, a)
    # raw C ABI with a pinned source and destination: the direct path
    lib = _lib.lib()
    src = torch.arange(6 << 20, dtype=torch.int32).pin_memory()
    dev = torch.empty_like(src, device="cuda")
    dst = torch.empty_like(src).pin_memory()
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(lib.svb_copy_h2d(dev.data_ptr(), src.data_ptr(), src.numel() * 4, st))
    _lib.check(lib.svb_copy_d2h(dst.data_ptr(), dev.data_ptr(), src.numel() * 4, st))
    assert torch.equal(dst, src)
    # through a public call: NumPy in, NumPy out
    img = rng.integers(0, 256, (1080, 1920, 3), dtype=np.uint8)
    g = svb.cvtColor(img, svb.COLOR_BGR2GRAY)
    assert isinstance(g, np.ndarray) and np.array_equal(g, oracle.bgr2gray(img))
