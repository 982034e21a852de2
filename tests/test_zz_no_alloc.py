"""GPU: svb_filter_speckles_s16_static -- the speckle filter for callers that cannot allocate device memory
(CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  It has no counterpart in the reference:
its scratch is a static device array inside libsvb.so.

Collected last on purpose.  Called through the C ABI and compared with the
existing entry point svb_filter_speckles_s16_dev (caller scratch) on two calls in a row -- the second with a
different shape and content, so anything the first call left in the shared array would show.  Integer output:
the tolerance is zero, as in the other speckle tests."""
import ctypes

import numpy as np
import pytest

from stereovision_b200 import _lib

pytestmark = pytest.mark.gpu

MAX_PIXELS = 2097152          # SVB_SPECKLE_STATIC_MAX_PIXELS (include/svb.h)


def _disparity_like(rng, B, H, W, inv):
    """Piecewise-constant int16 map with noise and invalid holes: components of every size."""
    d = np.zeros((B, H, W), np.int16)
    for b in range(B):
        base = rng.integers(0, 40, (H // 8 + 1, W // 8 + 1)) * 16
        d[b] = np.kron(base, np.ones((8, 8), np.int64))[:H, :W] + rng.integers(-3, 4, (H, W))
        speck = rng.random((H, W)) < 0.05
        d[b][speck] = rng.integers(0, 2000, int(speck.sum()))
        d[b][rng.random((H, W)) < 0.10] = inv
    return d


def test_static_scratch_speckle_filter_matches_the_allocating_entry_point():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    lib = _lib.lib()
    rng = np.random.default_rng(2718)
    inv, max_size, max_diff = -16, 60, 32
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    # two calls in a row: a large batch first, then a smaller, differently shaped one (stale labels / counts of the
    # first call lie beyond and underneath it), then the first shape again
    shapes = [(3, 240, 400), (1, 97, 131), (3, 240, 400)]
    free_before = None
    for k, (B, H, W) in enumerate(shapes):
        src = _disparity_like(rng, B, H, W, inv)
        want = torch.from_numpy(src).cuda()
        scratch = torch.empty(2 * want.numel(), dtype=torch.int32, device="cuda")
        _lib.check(lib.svb_filter_speckles_s16_dev(want.data_ptr(), B, H, W, inv, max_size, max_diff,
                                                   scratch.data_ptr(), st))
        got = torch.from_numpy(src).cuda()
        torch.cuda.synchronize()
        if k > 0:
            free_before = torch.cuda.mem_get_info()[0]
        _lib.check(lib.svb_filter_speckles_s16_static(got.data_ptr(), B, H, W, inv, max_size, max_diff, st))
        torch.cuda.synchronize()
        if k > 0:                                             # the call itself took no device memory (1 MiB of slack for
            assert torch.cuda.mem_get_info()[0] >= free_before - (1 << 20)   # whatever else shares the GPU)
        g, w = got.cpu().numpy(), want.cpu().numpy()
        assert int((w != src).sum()) > 0, "the filter removed nothing: the case does not test anything"
        assert np.array_equal(g, w), f"call {k}: {int((g != w).sum())} pixels differ from svb_filter_speckles_s16_dev"


def test_static_scratch_rejects_sizes_beyond_its_maximum():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    lib = _lib.lib()
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    H, W = 1024, 1024
    B = MAX_PIXELS // (H * W)                                 # exactly the maximum: accepted
    img = torch.full((B + 1, H, W), 7, dtype=torch.int16, device="cuda")
    assert lib.svb_filter_speckles_s16_static(img.data_ptr(), B, H, W, -16, 10, 16, st) == 0
    torch.cuda.synchronize()
    assert int((img != 7).sum()) == 0                         # one big component per image: kept
    rc = lib.svb_filter_speckles_s16_static(img.data_ptr(), B + 1, H, W, -16, 10, 16, st)
    assert rc == -2
    msg = lib.svb_last_error().decode()
    assert "exceeds" in msg and str(MAX_PIXELS) in msg
    with pytest.raises(NotImplementedError):
        _lib.check(rc)
    assert lib.svb_filter_speckles_s16_static(None, 1, 8, 8, -16, 10, 16, st) == -1


def test_static_scratch_call_can_be_captured_in_a_cuda_graph():
    """The use case in one test: a cudaMalloc / cudaFree / synchronising call inside a stream capture invalidates the
    capture, so an entry point that captures cleanly and replays correctly allocates nothing.  (The first call per
    device, which looks the scratch address up, is made before the capture, as include/svb.h says.)"""
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    lib = _lib.lib()
    rng = np.random.default_rng(99)
    B, H, W, inv = 2, 120, 200, -16
    src = _disparity_like(rng, B, H, W, inv)
    want = torch.from_numpy(src).cuda()
    scratch = torch.empty(2 * want.numel(), dtype=torch.int32, device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(lib.svb_filter_speckles_s16_dev(want.data_ptr(), B, H, W, inv, 60, 32, scratch.data_ptr(), st))
    warm = torch.from_numpy(src).cuda()
    _lib.check(lib.svb_filter_speckles_s16_static(warm.data_ptr(), B, H, W, inv, 60, 32, st))   # set-up outside the capture
    torch.cuda.synchronize()
    buf = torch.from_numpy(src).cuda()
    pristine = buf.clone()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        cs = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        rc = lib.svb_filter_speckles_s16_static(buf.data_ptr(), B, H, W, inv, 60, 32, cs)
    assert rc == 0, lib.svb_last_error()
    for _ in range(2):                                        # two replays on fresh input
        buf.copy_(pristine)
        graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(buf, want)
