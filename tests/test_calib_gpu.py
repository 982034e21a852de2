"""GPU: the calibration-file path end to end (BASELINE.json configs[0] 'synthetic calib .pkl'): a 7-key pickle on
disk -> svb.load_stereo_calibration / svb.load_stereo_calibration_with_scaling (host stereoRectify + the device map
kernel) -> maps bit-identical to what the reference's own loaders produced from the same file
(tests/golden/golden_calib.npz), Q to 1e-12 -> svb.remap consumes the dict as apply_stereo_rectification does."""
import os

import numpy as np
import pytest

import oracle
import stereovision_b200 as svb
from stereovision_b200.synth import make_pair, to_bgr, write_calibration_pkl
from tests.golden_cases import CALIB_CASE

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(ROOT, "tests", "golden", "golden_calib.npz"))


@pytest.fixture()
def pkl(tmp_path):
    p = tmp_path / "output" / "stereo_calibration_data.pkl"
    write_calibration_pkl(str(p), CALIB_CASE["W"], CALIB_CASE["H"])
    return str(p)


def _same(d, gold, tag):
    for k in ("left_map1", "left_map2", "right_map1", "right_map2"):
        m = d[k].cpu().numpy() if hasattr(d[k], "cpu") else d[k]
        assert np.array_equal(m, gold[f"{tag}_{k}"]), (tag, k, int((m != gold[f"{tag}_{k}"]).sum()))
    assert np.allclose(d["Q"], gold[f"{tag}_Q"], rtol=1e-12, atol=1e-12)
    assert np.array_equal(np.array([d["roi1"], d["roi2"]]), gold[f"{tag}_roi"])
    assert d["baseline"] == float(gold[f"{tag}_baseline"])


def test_load_stereo_calibration_from_pkl(gold, pkl, monkeypatch):
    d = svb.load_stereo_calibration(pkl)
    _same(d, gold, "plain")
    assert tuple(d["img_size"]) == tuple(gold["plain_size"])
    # the reference's relative default path (depth_map.py:22)
    monkeypatch.chdir(os.path.dirname(os.path.dirname(pkl)))
    _same(svb.load_stereo_calibration(), gold, "plain")
    monkeypatch.chdir(os.path.dirname(pkl))
    assert svb.load_stereo_calibration() is None


@pytest.mark.parametrize("scale", CALIB_CASE["scales"])
def test_load_stereo_calibration_with_scaling(gold, pkl, scale):
    tag = f"scale{int(round(scale * 100))}"
    d = svb.load_stereo_calibration_with_scaling(scale, pkl)
    _same(d, gold, tag)
    assert tuple(d["img_size_proc"]) == tuple(gold[f"{tag}_size"]) and d["focal_length"] == float(gold[f"{tag}_focal_length"])
    assert d["img_size_orig"] == (CALIB_CASE["W"], CALIB_CASE["H"]) and d["scale_factor"] == scale
    dd = svb.load_stereo_calibration_with_scaling(scale, pkl, device_maps=True)
    assert dd["left_map1"].is_cuda
    _same(dd, gold, tag)


def test_frame_path_from_the_pkl(gold, pkl):
    """fused_depth_map.apply_stereo_rectification (:444-500) from the file: resize both frames to img_size_proc
    (every processed frame at PROCESSING_SCALE 0.33), then remap with the loaded maps -- host maps and device maps
    give the same bytes, equal to the oracle fed with the reference's golden maps."""
    import torch
    scale = 0.33
    d = svb.load_stereo_calibration_with_scaling(scale, pkl)
    dd = svb.load_stereo_calibration_with_scaling(scale, pkl, device_maps=True)
    l, r, _ = make_pair(CALIB_CASE["H"], CALIB_CASE["W"], 32, seed=77)
    lb, rb = to_bgr(l), to_bgr(r)
    size = d["img_size_proc"]
    tag = "scale33"
    for img, side in ((lb, "left"), (rb, "right")):
        small = svb.resize(img, size, interpolation=svb.INTER_LINEAR)
        assert np.array_equal(small, oracle.resize(img, size))
        rect = svb.remap(small, d[f"{side}_map1"], d[f"{side}_map2"], svb.INTER_LINEAR)
        want = oracle.remap(small, gold[f"{tag}_{side}_map1"], gold[f"{tag}_{side}_map2"])
        assert np.array_equal(rect, want)
        rect_dev = svb.remap(svb.resize(torch.from_numpy(img).cuda(), size), dd[f"{side}_map1"], dd[f"{side}_map2"])
        assert np.array_equal(rect_dev.cpu().numpy(), want)
