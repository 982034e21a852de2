"""stereovision_b200 -- B200-native (sm_100a) engine for the per-frame hot path of
AlexGr5/StereoVision (depth_map.py / fused_depth_map.py), behind cv2's call surface.

Hand-written CUDA behind a C ABI (include/svb.h, libsvb.so); Python + torch are plumbing.
No CPU fallback: without the built library or a CUDA device, calls raise.
"""
from .cv_api import (  # noqa: F401
    STEREO_SGBM_MODE_SGBM, STEREO_SGBM_MODE_HH, STEREO_SGBM_MODE_SGBM_3WAY, STEREO_SGBM_MODE_HH4,
    StereoSGBM_MODE_SGBM, StereoSGBM_MODE_HH, StereoSGBM_MODE_SGBM_3WAY, StereoSGBM_MODE_HH4,
    STEREO_BM_PREFILTER_NORMALIZED_RESPONSE, STEREO_BM_PREFILTER_XSOBEL,
    INTER_LINEAR, COLOR_BGR2GRAY,
    StereoSGBM, StereoBM, StereoSGBM_create, StereoBM_create,
    remap, resize, cvtColor, medianBlur, filterSpeckles, validateDisparity, reprojectImageTo3D, disparity_to_depth,
    initUndistortRectifyMap, CV_16SC2,
    release_workspaces, pool_size, invalidate_map_cache,
)

from .fusion import (  # noqa: F401
    normalize_disparity, calibrate_midas_to_stereo, normalize_to_stereo_range, fuse_depth_maps,
    detect_camera_occlusion, occlusion_stats, percentiles, OcclusionMonitor,
)

from .mono import (  # noqa: F401
    bilateralFilter, midas_postprocess, upsample_normalize, contrast_confidence, flow_residual_depth,
)

from .display import (  # noqa: F401
    COLORMAP_JET, COLORMAP_TURBO, applyColorMap, clip_colormap, depth_colormap, depth_display_colormap,
    disparity_colormap, legend_range, legend_depth_range,
)
from .calib import (  # noqa: F401
    stereoRectify, Rodrigues, load_stereo_calibration, load_stereo_calibration_with_scaling, CALIB_ZERO_DISPARITY,
)
from . import frame  # noqa: F401  (frame.apply_stereo_rectification / frame.create_depth_map: one upload, one download)
from ._lib import SvbError  # noqa: F401

__version__ = "0.1.0"
