"""CPU: the oracle at the BASELINE.json sizes against the committed SHA-256 of cv2's own output
(tests/golden/fullsize_sha256.json, made by tests/golden/make_golden_fullsize.py with cv2 4.13.0).
This pins the oracle -- and through it the CUDA path -- at the graded shapes, not just at thumbnails:
1080p has 60 bands x 1664 cost columns per image, 3WAY at D=320 saturates S."""
import numpy as np
import pytest

import oracle
from stereovision_b200.synth import make_pair
from tests.fullsize_cases import FULLSIZE_CASES, load_hashes, sha256_of


@pytest.mark.parametrize("name,seed", [
    ("c3_720p_d128_sgbm", 1),            # BASELINE config 3: about 6 s
    ("script_480p_d320_3way", 1),        # depth_map.py's own parameters
    ("c1_480p_d64_3way", 1),
    ("c1_480p_d64_sgbm", 1),
    ("c4_1080p_d256_hh", 1),             # BASELINE config 4, the headline: about 26 s, 5.5 GB of host memory
])
def test_oracle_matches_cv2_hash_at_full_size(name, seed):
    c = FULLSIZE_CASES[name]
    want = load_hashes()["tests"][f"{name}/seed{seed}"]
    l, r, _ = make_pair(c["H"], c["W"], c["prm"]["numDisparities"], seed=seed)
    got = oracle.sgbm_compute(l, r, **c["prm"])
    assert int((got >= 0).sum()) == want["valid"] and int(got.astype(np.int64).sum()) == want["sum"]
    assert sha256_of(got) == want["sha256"]


def test_bench_workloads_agree_with_the_pinned_table():
    """bench.py hashes its own outputs against the 'bench' group: the tables must describe the same work."""
    import bench
    from tests.fullsize_cases import BENCH_WORKLOADS, bench_params
    for name, wl in BENCH_WORKLOADS.items():
        b = bench.WORKLOADS[name]
        assert (b["H"], b["W"], b["D"], b["mode"]) == (wl["H"], wl["W"], wl["D"], wl["mode"])
        assert bench.sgbm_params(b) == bench_params(wl)
    h = load_hashes()
    assert h["cv2_version"].startswith("4.13")
    for r in range(8):
        for i in range(4):
            assert f"sgbm_1080p_d256_hh/seed{1000 * r + i}" in h["bench"]
