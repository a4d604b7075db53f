"""Host-side mirror of /root/reference/pcdet/ops/roiaware_pool3d/roiaware_pool3d_utils.py for the
hot-path functions (same names, argument meaning and return types)."""
import torch

from ...utils import common_utils
from . import roiaware_pool3d_cuda


def points_in_boxes_cpu(points, boxes):
    """reference :9-25.  points (N,3), boxes (M,7) numpy or CPU tensors -> (M,N) int32 0/1 mask
    (numpy in -> numpy out).  Margin 1e-2; computed on the GPU through host buffers."""
    assert boxes.shape[1] == 7
    assert points.shape[1] == 3
    points, is_numpy = common_utils.check_numpy_to_torch(points)
    boxes, is_numpy = common_utils.check_numpy_to_torch(boxes)
    mask = torch.zeros((boxes.shape[0], points.shape[0]), dtype=torch.int32)
    roiaware_pool3d_cuda.points_in_boxes_cpu(boxes.float().contiguous(), points.float().contiguous(), mask)
    return mask.numpy() if is_numpy else mask


def points_in_boxes_gpu(points, boxes):
    """reference :28-41.  points (B,N,3), boxes (B,M,7) CUDA -> (B,N) int32 first containing box or -1."""
    assert boxes.shape[0] == points.shape[0]
    assert boxes.shape[2] == 7 and points.shape[2] == 3
    B, N, _ = points.shape
    box_idxs_of_pts = torch.full((B, N), -1, dtype=torch.int32, device=points.device)
    roiaware_pool3d_cuda.points_in_boxes_gpu(boxes.contiguous(), points.contiguous(), box_idxs_of_pts)
    return box_idxs_of_pts
