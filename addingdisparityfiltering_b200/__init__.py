"""B200-native drop-in for the ximgproc disparity-filtering hot path
(cv::ximgproc::DisparityWLSFilter::filter and the FastGlobalSmootherFilter it drives).

CUDA kernels for sm_100a behind a C-ABI (include/dwls_b200.h); this package is the Python
mirror of the reference's CV_WRAP API on numpy arrays plus the batch / multi-GPU entry points.
"""
from ._lib import Context, DwlsError, build, default_context, lib  # noqa: F401
from .ximgproc import (DisparityWLSFilter, FastGlobalSmootherFilter, computeBadPixelPercent, computeMSE,  # noqa: F401
                       createDisparityWLSFilter, createDisparityWLSFilterGeneric, createFastGlobalSmootherFilter,
                       createRightMatcher, fastGlobalSmootherFilter, fgs_weights, getDisparityVis, readGT)
from .stereo import StereoBM  # noqa: F401,E402
