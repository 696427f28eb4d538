"""foundationstereo_b200 -- B200 (sm_100a) kernels for FoundationStereo's cost-volume and
iterative-refinement hot path, behind the reference's own Python interface.

    from foundationstereo_b200 import install
    install(model)        # core/foundation_stereo.py then runs unchanged on the new path

The arithmetic lives in libfsb200.so (include/fsb200.h, csrc/*.cu).  There is no CPU or
eager fallback: importing the ops without the built library raises.
"""
from ._lib import FsbError, launch_count, lib  # noqa: F401
from .submodule import build_gwc_volume, build_concat_volume, build_combined_volume  # noqa: F401
from .submodule import (disparity_regression, softmax_disparity_regression, context_upsample,  # noqa: F401
                        softmax_context_upsample)
from .geometry import Combined_Geo_Encoding_Volume  # noqa: F401
from .update import SelectiveUpdateBlock  # noqa: F401
from .hotpath import StereoHotPath  # noqa: F401
from . import shard  # noqa: F401
from .patch import install, install_update_block, uninstall  # noqa: F401

__all__ = [
    "build_gwc_volume", "build_concat_volume", "build_combined_volume", "Combined_Geo_Encoding_Volume",
    "SelectiveUpdateBlock", "StereoHotPath", "install", "install_update_block", "uninstall", "FsbError",
    "launch_count", "lib", "disparity_regression", "softmax_disparity_regression", "context_upsample",
    "softmax_context_upsample",
]
