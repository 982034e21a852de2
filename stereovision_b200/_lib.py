"""ctypes binding of libsvb.so (the C ABI declared in include/svb.h).

The library is built in-tree by ``stereovision_b200/csrc/Makefile`` (nvcc, sm_100a).
There is no CPU fallback: if the library is missing, or no CUDA device is present,
every compute call raises.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.environ.get("SVB_LIBRARY") or os.path.join(_HERE, "libsvb.so")   # SVB_LIBRARY: developer override (experiments)
_lib = None

c_int, c_size_t, c_void_p, c_double = ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_double


class SgbmParams(ctypes.Structure):
    _fields_ = [(n, c_int) for n in (
        "minDisparity", "numDisparities", "blockSize", "P1", "P2", "disp12MaxDiff",
        "preFilterCap", "uniquenessRatio", "speckleWindowSize", "speckleRange", "mode")]


class BmParams(ctypes.Structure):
    _fields_ = [(n, c_int) for n in (
        "minDisparity", "numDisparities", "blockSize", "preFilterCap", "textureThreshold",
        "uniquenessRatio", "speckleWindowSize", "speckleRange", "disp12MaxDiff", "preFilterType", "preFilterSize")]


class FuseParams(ctypes.Structure):
    _fields_ = [(n, ctypes.c_double) for n in (
        "stereo_weight", "midas_fill_weight", "flow_fill_weight", "stereo_conf_threshold", "flow_hole_threshold",
        "bilateral_sigma_color", "bilateral_sigma_space")] + [(n, ctypes.c_int) for n in (
        "blend_ksize", "bilateral_d", "use_midas_fill", "use_flow_fill")]


# name -> (restype, argtypes); every symbol include/svb.h declares
SIGNATURES = {
    "svb_last_error": (ctypes.c_char_p, []),
    "svb_version": (c_int, []),
    "svb_init": (c_int, [c_int]),
    "svb_device_count": (c_int, []),
    "svb_copy_h2d": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p]),
    "svb_copy_d2h": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p]),
    "svb_sgbm_create": (c_int, [ctypes.POINTER(SgbmParams), ctypes.POINTER(c_void_p)]),
    "svb_sgbm_destroy": (None, [c_void_p]),
    "svb_sgbm_workspace_bytes": (c_size_t, [c_void_p, c_int, c_int, c_int]),
    "svb_sgbm_set_workspace_limit": (c_int, [c_void_p, c_size_t]),
    "svb_sgbm_compute_dev": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "svb_sgbm_compute_host": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "svb_sgbm_compute_cn_dev": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "svb_sgbm_compute_cn_host": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "svb_sgbm_compute_pitched_dev": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, ctypes.c_ssize_t,
                                             ctypes.c_ssize_t, c_void_p, c_void_p]),
    "svb_sgbm_set_debug": (c_int, [c_void_p, c_int]),
    "svb_sgbm_debug_stages": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    "svb_sgbm_last_launches": (c_int, [c_void_p]),
    "svb_sgbm_check": (c_int, [c_void_p]),
    "svb_sgbm_debug_trip_error": (c_int, [c_void_p]),
    "svb_sgbm_debug_force_wave": (c_int, [c_void_p, c_int]),
    "svb_sgbm_set_profiling": (c_int, [c_void_p, c_int]),
    "svb_sgbm_get_profile": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "svb_bm_create": (c_int, [ctypes.POINTER(BmParams), ctypes.POINTER(c_void_p)]),
    "svb_bm_destroy": (None, [c_void_p]),
    "svb_bm_compute_dev": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "svb_bm_compute_host": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "svb_bm_last_launches": (c_int, [c_void_p]),
    "svb_validate_disparity_dev": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                           c_void_p, c_void_p]),
    "svb_remap_u8_dev": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int,
                                 c_void_p, c_int, c_void_p]),
    "svb_bgr2gray_dev": (c_int, [c_void_p, c_size_t, c_void_p, c_void_p]),
    "svb_resize_u8_dev": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int, c_int, c_void_p]),
    "svb_resize_f32_dev": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_int, c_int, c_void_p]),
    "svb_median3_s16_dev": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "svb_median5_f32_dev": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "svb_filter_speckles_s16_dev": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "svb_filter_speckles_s16_static": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "svb_disp_to_depth_dev": (c_int, [c_void_p, c_int, c_int, c_int, c_double, c_double, c_double, c_double,
                                      c_double, c_void_p, c_void_p, c_void_p]),
    "svb_reproject3d_dev": (c_int, [c_void_p, c_int, c_int, c_int, c_int, ctypes.POINTER(c_double), c_int,
                                    c_void_p, c_void_p, c_void_p]),
    "svb_init_undistort_rectify_map_dev": (c_int, [ctypes.POINTER(c_double), ctypes.POINTER(c_double), c_int,
                                                   ctypes.POINTER(c_double), ctypes.POINTER(c_double), c_int, c_int,
                                                   c_int, c_void_p, c_void_p, c_void_p]),
    "svb_normalize_disparity_dev": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "svb_range_scratch_bytes": (c_size_t, [c_int]),
    "svb_calibrate_midas_dev": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p,
                                        c_size_t, c_void_p]),
    "svb_normalize_to_stereo_range_dev": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p,
                                                  c_size_t, c_void_p]),
    "svb_percentiles_f32_dev": (c_int, [c_void_p, c_void_p, ctypes.c_float, c_int, c_int, c_size_t, c_void_p,
                                        c_void_p, c_void_p, c_size_t, c_void_p]),
    "svb_fuse_scratch_bytes": (c_size_t, [c_int, c_int, c_int]),
    "svb_fuse_depth_maps_dev": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                        ctypes.POINTER(FuseParams), c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "svb_occlusion_scratch_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "svb_occlusion_stats_u8_dev": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_size_t,
                                           c_void_p]),
    "svb_bilateral_scratch_bytes": (c_size_t, [c_int]),
    "svb_bilateral_f32_dev": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_double, c_double, c_void_p, c_void_p,
                                      c_size_t, c_void_p]),
    "svb_midas_scratch_bytes": (c_size_t, [c_int]),
    "svb_midas_upsample_normalize_dev": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_int, c_int, c_void_p,
                                                 c_size_t, c_void_p]),
    "svb_contrast_confidence_u8_dev": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_size_t, c_void_p]),
    "svb_flow_residual_depth_dev": (c_int, [c_void_p, c_int, c_int, c_int, ctypes.POINTER(c_double), c_void_p,
                                            c_void_p, c_int, c_double, c_void_p]),
    "svb_disp_to_depth_f64_dev": (c_int, [c_void_p, c_int, c_int, c_int, c_double, c_double, c_double, c_double,
                                          c_double, c_void_p, c_void_p, c_void_p]),
    "svb_display_scratch_bytes": (c_size_t, [c_int]),
    "svb_colormap_dev": (c_int, [c_void_p, c_int, c_int, c_size_t, c_int, ctypes.POINTER(c_double), c_void_p,
                                 c_void_p, c_void_p, c_size_t, c_void_p]),
    "svb_range_nonzero_u8_dev": (c_int, [c_void_p, c_int, c_size_t, c_void_p, c_void_p, c_size_t, c_void_p]),
}


def build(force=False):
    """Compile csrc/*.cu -> libsvb.so with nvcc for sm_100a (cross-compiles without a GPU)."""
    args = ["make", "-C", os.path.join(_HERE, "csrc"), "-j8"]
    if force:
        args.append("-B")
    subprocess.check_call(args, stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise RuntimeError(
                f"{_SO} is missing: build it with `make -C stereovision_b200/csrc` "
                "(python -c 'import __graft_entry__ as g; g.build()'). There is no CPU fallback.")
        l = ctypes.CDLL(_SO)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


class SvbError(RuntimeError):
    pass


def check(rc):
    """Map svb_status to Python exceptions (cv2 raises cv2.error; the reference catches Exception)."""
    if rc == 0:
        return
    msg = lib().svb_last_error().decode("utf-8", "replace")
    if rc == -1:
        raise ValueError(msg)
    if rc == -2:
        raise NotImplementedError(msg)
    if rc == -4:
        raise MemoryError(msg)
    raise SvbError(msg)
