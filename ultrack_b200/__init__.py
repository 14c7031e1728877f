"""ultrack_b200 -- B200-native candidate generation for ultrack: `segment()` and `link()`.

Drop-in for `ultrack.segment` / `ultrack.link` (ultrack/__init__.py:22,24) on top of a C-ABI
library of hand-written sm_100a CUDA kernels (include/ultrack_b200.h).  No CPU fallback.
"""
from ultrack_b200.config import (  # noqa: F401
    DataConfig,
    LinkingConfig,
    MainConfig,
    SegmentationConfig,
    load_config,
    save_config,
)
from ultrack_b200.edge import labels_to_contours  # noqa: F401
from ultrack_b200.flow import add_flow  # noqa: F401
from ultrack_b200.link import link  # noqa: F401
from ultrack_b200.segment import segment  # noqa: F401

__all__ = ["segment", "link", "add_flow", "labels_to_contours", "MainConfig", "DataConfig", "SegmentationConfig",
           "LinkingConfig", "load_config", "save_config"]
