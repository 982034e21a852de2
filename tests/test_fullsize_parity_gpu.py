"""GPU: bit-exact parity AT THE GRADED SIZES.  Every BASELINE.json matcher configuration (config 3
720p/D128/MODE_SGBM, config 4 1080p/D256/MODE_HH, the script's own 640x480/D320/bs7/3WAY, config 1) is run
through the cv2-shaped API and through the raw C ABI -- host path and device path, a batch that spans the
host pipeline's four sub-ranges twice, and a workspace limit that forces two chunks -- and compared with the
committed SHA-256 of cv2's own output (tests/golden/fullsize_sha256.json) and, where it finishes in seconds, with
the oracle itself."""
import ctypes

import numpy as np
import pytest

import oracle
import stereovision_b200 as svb
from stereovision_b200 import _lib
from stereovision_b200.synth import make_pair
from tests.fullsize_cases import FULLSIZE_CASES, load_hashes, sha256_of

pytestmark = pytest.mark.gpu

PARAM_ORDER = ("minDisparity", "numDisparities", "blockSize", "P1", "P2", "disp12MaxDiff", "preFilterCap",
               "uniquenessRatio", "speckleWindowSize", "speckleRange", "mode")


def _torch():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch


def _explain(got, l, r, prm, want):
    """A hash says nothing about WHERE: on a mismatch run the oracle (it equals cv2 at these sizes,
    tests/test_fullsize_oracle.py) and report the differing pixels."""
    ref = oracle.sgbm_compute(l, r, **prm)
    bad = np.argwhere(got != ref)
    return (f"sha256 {sha256_of(got)[:16]} != cv2 {want['sha256'][:16]}; {len(bad)} pixels differ from the oracle, "
            f"first at {bad[:5].tolist()}; oracle itself {'==' if sha256_of(ref) == want['sha256'] else '!='} cv2")


@pytest.mark.parametrize("name", sorted(FULLSIZE_CASES))
def test_single_pair_matches_cv2_hash(name):
    _torch()
    c = FULLSIZE_CASES[name]
    hashes = load_hashes()["tests"]
    m = svb.StereoSGBM_create(**c["prm"])
    for seed in c["seeds"]:
        want = hashes[f"{name}/seed{seed}"]
        l, r, _ = make_pair(c["H"], c["W"], c["prm"]["numDisparities"], seed=seed)
        got = m.compute(l, r)                                 # NumPy in / NumPy out, the reference's call (depth_map.py:909)
        assert got.dtype == np.int16 and got.shape == l.shape
        assert sha256_of(got) == want["sha256"], _explain(got, l, r, c["prm"], want)


@pytest.mark.parametrize("name", ["c3_720p_d128_sgbm", "script_480p_d320_3way", "c1_480p_d64_sgbm"])
def test_single_pair_equals_oracle_array(name):
    """The same against the oracle's array (not only its hash) where the oracle takes seconds."""
    _torch()
    c = FULLSIZE_CASES[name]
    l, r, _ = make_pair(c["H"], c["W"], c["prm"]["numDisparities"], seed=c["seeds"][0])
    got = svb.StereoSGBM_create(**c["prm"]).compute(l, r)
    ref = oracle.sgbm_compute(l, r, **c["prm"])
    assert int((got != ref).sum()) == 0


@pytest.mark.parametrize("name,B", [("c4_1080p_d256_hh", 10), ("c3_720p_d128_sgbm", 12), ("script_480p_d320_3way", 9)])
def test_batches_host_pipeline_chunks_and_device_path(name, B):
    """B >= 2 x SVB_HOST_PIPE pairs (two distinct scenes alternating, plus a third at the end so that the last
    sub-range is ragged): (a) host buffers through the C ABI with the default workspace -- one chunk, four pipelined
    sub-ranges; (b) the same with a workspace limit that forces two chunks; (c) device buffers on the current
    torch stream.  Every output image must hash to cv2's."""
    torch = _torch()
    c = FULLSIZE_CASES[name]
    H, W, prm = c["H"], c["W"], c["prm"]
    hashes = load_hashes()["tests"]
    seeds = [c["seeds"][i % 2] for i in range(B - 1)] + [c["seeds"][0]]
    scenes = {s: make_pair(H, W, prm["numDisparities"], seed=s)[:2] for s in set(seeds)}
    L = np.stack([scenes[s][0] for s in seeds])
    R = np.stack([scenes[s][1] for s in seeds])
    lib = _lib.lib()
    h = ctypes.c_void_p()
    _lib.check(lib.svb_sgbm_create(ctypes.byref(_lib.SgbmParams(*[prm[k] for k in PARAM_ORDER])), ctypes.byref(h)))
    try:
        def check(out, what):
            for i, s in enumerate(seeds):
                want = hashes[f"{name}/seed{s}"]["sha256"]
                assert sha256_of(out[i]) == want, f"{what}: image {i} (seed {s}) differs from cv2"

        out = np.empty((B, H, W), np.int16)
        _lib.check(lib.svb_sgbm_compute_host(h, L.ctypes.data, R.ctypes.data, B, H, W, out.ctypes.data))
        check(out, "host path, one chunk")

        per_pair = lib.svb_sgbm_workspace_bytes(h, 1, H, W)
        lib.svb_sgbm_set_workspace_limit(h, ctypes.c_size_t(per_pair * ((B + 1) // 2) + (1 << 20)))
        assert lib.svb_sgbm_workspace_bytes(h, B, H, W) < per_pair * B          # really fewer pairs than B per chunk
        out2 = np.empty((B, H, W), np.int16)
        _lib.check(lib.svb_sgbm_compute_host(h, L.ctypes.data, R.ctypes.data, B, H, W, out2.ctypes.data))
        check(out2, "host path, two chunks")
        lib.svb_sgbm_set_workspace_limit(h, ctypes.c_size_t(0))

        Ld, Rd = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
        Od = torch.empty((B, H, W), dtype=torch.int16, device="cuda")
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        _lib.check(lib.svb_sgbm_compute_dev(h, Ld.data_ptr(), Rd.data_ptr(), B, H, W, Od.data_ptr(), st))
        _lib.check(lib.svb_sgbm_check(h))
        check(Od.cpu().numpy(), "device path")
    finally:
        lib.svb_sgbm_destroy(h)
