"""stereo_vision_b200 — B200-native (sm_100a) stereo depth engine behind the cv2 call surface of the
jacksonroah/stereo-vision hot path: rectify -> StereoSGBM -> reprojectImageTo3D -> per-box depth."""
from .api import (  # noqa: F401
    BORDER_CONSTANT, CV_16SC2, CV_32FC1, INTER_LINEAR, STEREO_SGBM_MODE_HH, STEREO_SGBM_MODE_HH4,
    STEREO_SGBM_MODE_SGBM, STEREO_SGBM_MODE_SGBM_3WAY, StereoSGBM, StereoSGBM_create, box_median_depth,
    filterSpeckles, initUndistortRectifyMap, last_stage_ms, medianBlur, packed_alu_peak, release_workspaces, remap,
    reprojectImageTo3D, set_latency_form, set_stage_timing,
    COLOR_BGR2GRAY, cvtColor, triangulatePoints, triangulate_stereo, undistort, undistortPoints,
)
from ._lib import SvbError  # noqa: F401
from .calib import StereoRig, load_calibration  # noqa: F401
from .stream import HostPipeline  # noqa: F401

__version__ = "0.1.0"
