"""Frame-level mirror of the reference's two per-frame functions, one upload and one download per frame:

    depth_map.apply_stereo_rectification(left_img, right_img, stereo_calib)          depth_map.py:779-834
                                                     (fused_depth_map.py:444-500 is the same with 'img_size_proc')
    depth_map.create_depth_map(left_img, right_img, stereo_calib, min_depth, max_depth)   depth_map.py:837-946

Calling the cv2-shaped functions one by one with NumPy arrays (the literal `cv2.` -> `svb.` substitution of
INTEGRATION.md) copies every intermediate image to the host and back.  These two keep the whole chain on the device:
the frames go up once, resize / remap / gray / SGBM / median / depth / colour map run back to back on the current
stream, and only what the reference returns comes down.  Same arguments, same return values, same error behaviour
(every exception is caught and zeros come back, depth_map.py:941-946); the script's module globals MIN_DISP, NUM_DISP,
WINDOW_SIZE are keyword arguments with the script's values.
"""
import numpy as np

from . import cv_api as _cv
from . import display as _display
from .cv_api import _is_tensor, _torch


def _up(img):
    """NumPy -> CUDA tensor (one copy); tensors pass through."""
    if _is_tensor(img):
        return img
    return _cv._upload(np.ascontiguousarray(img))


def _down(t, as_tensor):
    return t if as_tensor else _cv._download(t)


def apply_stereo_rectification(left_img, right_img, stereo_calib, as_tensor=False):
    """Rectified (left, right).  `stereo_calib` is the dict of svb.load_stereo_calibration[_with_scaling] (maps as NumPy
    arrays or, with device_maps=True, as CUDA tensors: no per-frame map traffic at all).  Frames of another size are
    resized to the calibration / processing size first, as the reference does."""
    if stereo_calib is None:
        return left_img, right_img
    try:
        size = stereo_calib.get("img_size_proc", stereo_calib.get("img_size"))
        out = []
        for img, side in ((left_img, "left"), (right_img, "right")):
            t = _up(img)
            h, w = int(t.shape[0]), int(t.shape[1])
            if size is not None and (w, h) != (int(size[0]), int(size[1])):
                t = _cv.resize(t, (int(size[0]), int(size[1])), interpolation=_cv.INTER_LINEAR)
            out.append(_cv.remap(t, stereo_calib[f"{side}_map1"], stereo_calib[f"{side}_map2"], _cv.INTER_LINEAR))
        return _down(out[0], as_tensor), _down(out[1], as_tensor)
    except Exception as e:                                    # depth_map.py:832-834: report, hand the inputs back
        import warnings
        warnings.warn(f"apply_stereo_rectification: {e}")
        return left_img, right_img


def create_depth_map(left_img, right_img, stereo_calib=None, min_depth=0.3, max_depth=2.0, *, min_disp=0, num_disp=320,
                     window_size=7, as_tensor=False):
    """(depth_final, disparity float32, depth_colormap BGR uint8) like depth_map.create_depth_map.  depth_final is
    float32 with the script's fx = 700 and float64 when fx comes from stereo_calib['calibration_data']['camera_matrix']
    (a NumPy float64: NEP 50 promotes the reference's whole expression)."""
    try:
        l, r = _up(left_img), _up(right_img)
        if tuple(l.shape[:2]) != tuple(r.shape[:2]):          # ensure_same_size (depth_map.py:39-70): the smaller of the two
            h, w = min(int(l.shape[0]), int(r.shape[0])), min(int(l.shape[1]), int(r.shape[1]))
            l, r = _cv.resize(l, (w, h)), _cv.resize(r, (w, h))
        gl = _cv.cvtColor(l, _cv.COLOR_BGR2GRAY) if l.ndim == 3 else l
        gr = _cv.cvtColor(r, _cv.COLOR_BGR2GRAY) if r.ndim == 3 else r
        ws = int(window_size)
        stereo = _cv.StereoSGBM_create(minDisparity=int(min_disp), numDisparities=int(num_disp), blockSize=ws,
                                       P1=8 * 3 * ws ** 2, P2=32 * 3 * ws ** 2, disp12MaxDiff=1, uniquenessRatio=10,
                                       speckleWindowSize=100, speckleRange=32, preFilterCap=63,
                                       mode=_cv.STEREO_SGBM_MODE_SGBM_3WAY)
        d16 = stereo.compute(gl, gr)
        if stereo_calib is not None and "calibration_data" in stereo_calib:
            fx = stereo_calib["calibration_data"]["camera_matrix"][0, 0]
        else:
            fx = 700
        baseline = 0.08
        depth, disparity = _cv.disparity_to_depth(d16, fx=fx, baseline=baseline, min_disp=min_disp, min_depth=min_depth,
                                                  max_depth=max_depth)
        cmap = _display.depth_colormap(disparity, fx=fx, baseline=baseline, min_depth=min_depth, max_depth=max_depth)
        if not as_tensor:
            stereo.check()
        return _down(depth, as_tensor), _down(disparity, as_tensor), _down(cmap, as_tensor)
    except Exception as e:                                    # depth_map.py:941-946: zeros and the colour map of zeros
        import warnings
        warnings.warn(f"create_depth_map: {e}")
        h, w = int(left_img.shape[0]), int(left_img.shape[1])
        zeros = np.zeros((h, w), np.float32)
        empty = _display.applyColorMap(np.zeros((h, w), np.uint8), _display.COLORMAP_TURBO)
        return zeros, zeros.copy(), empty
