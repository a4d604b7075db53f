"""multimodal_gar_b200 - B200-native (sm_100a) LiDAR hot path of MGAR-net.

Drop-in for the reference's pcdet.ops surface on this path (points_in_boxes, roipoint_pool3d,
pointnet2_stack FPS / ball query / grouping / SA shared MLP) over a C-ABI CUDA library.
"""
from . import _lib  # noqa: F401

__all__ = ["_lib"]
