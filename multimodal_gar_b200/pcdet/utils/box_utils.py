"""Box helpers of the reference's pcdet/utils/box_utils.py used on the LiDAR hot path."""
import torch

from . import common_utils


def enlarge_box3d(boxes3d, extra_width=(0, 0, 0)):
    """pcdet/utils/box_utils.py:187-201: dx,dy,dz += extra_width.  boxes3d (N,7).

    The reference indexes `new_tensor(extra_width)[None, :]` and therefore crashes on a scalar
    (SURVEY.md D5); a scalar is accepted here and broadcast to the three extents."""
    boxes3d, is_numpy = common_utils.check_numpy_to_torch(boxes3d)
    large = boxes3d.clone()
    extra = boxes3d.new_tensor(extra_width).reshape(-1)
    if extra.numel() == 1:
        extra = extra.expand(3)
    large[:, 3:6] += extra[None, :]
    return large.numpy() if is_numpy else large


def remove_points_in_boxes3d(points, boxes3d):
    """pcdet/utils/box_utils.py:117-131: drop the points that fall in any box (CPU-mask semantics)."""
    from ..ops.roiaware_pool3d import roiaware_pool3d_utils
    boxes3d, is_numpy = common_utils.check_numpy_to_torch(boxes3d)
    points, is_numpy = common_utils.check_numpy_to_torch(points)
    mask = roiaware_pool3d_utils.points_in_boxes_cpu(points[:, 0:3], boxes3d)
    points = points[mask.sum(dim=0) == 0]
    return points.numpy() if is_numpy else points
