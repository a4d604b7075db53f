"""Host-side mirror of /root/reference/pcdet/ops/roiaware_pool3d/roiaware_pool3d_utils.py for the
hot-path functions (same names, argument meaning and return types)."""
import torch
import torch.nn as nn
from torch.autograd import Function

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


class RoIAwarePool3d(nn.Module):
    """reference :43-51.  forward(rois (N,7), pts (npoints,3), pts_feature (npoints,C), pool_method 'max'|'avg')
    -> (N, out_x, out_y, out_z, C)."""

    def __init__(self, out_size, max_pts_each_voxel=128):
        super().__init__()
        self.out_size = out_size
        self.max_pts_each_voxel = max_pts_each_voxel

    def forward(self, rois, pts, pts_feature, pool_method='max'):
        assert pool_method in ['max', 'avg']
        return RoIAwarePool3dFunction.apply(rois, pts, pts_feature, self.out_size, self.max_pts_each_voxel, pool_method)


class RoIAwarePool3dFunction(Function):
    """reference :54-107: the gradient flows to pts_feature only."""

    @staticmethod
    def forward(ctx, rois, pts, pts_feature, out_size, max_pts_each_voxel, pool_method):
        assert rois.shape[1] == 7 and pts.shape[1] == 3
        if isinstance(out_size, int):
            out_x = out_y = out_z = out_size
        else:
            assert len(out_size) == 3 and all(isinstance(k, int) for k in out_size)
            out_x, out_y, out_z = out_size
        num_rois, num_channels, num_pts = rois.shape[0], pts_feature.shape[-1], pts.shape[0]
        shape = (num_rois, out_x, out_y, out_z)
        pooled_features = pts_feature.new_zeros(shape + (num_channels,))
        argmax = pts_feature.new_zeros(shape + (num_channels,), dtype=torch.int)
        pts_idx_of_voxels = pts_feature.new_zeros(shape + (max_pts_each_voxel,), dtype=torch.int)
        method = {'max': 0, 'avg': 1}[pool_method]
        roiaware_pool3d_cuda.forward(rois.contiguous(), pts.contiguous(), pts_feature.contiguous(), argmax,
                                     pts_idx_of_voxels, pooled_features, method)
        ctx.roiaware_pool3d_for_backward = (pts_idx_of_voxels, argmax, method, num_pts, num_channels)
        return pooled_features

    @staticmethod
    def backward(ctx, grad_out):
        pts_idx_of_voxels, argmax, method, num_pts, num_channels = ctx.roiaware_pool3d_for_backward
        grad_in = grad_out.new_zeros((num_pts, num_channels))
        roiaware_pool3d_cuda.backward(pts_idx_of_voxels, argmax, grad_out.contiguous(), grad_in, method)
        return None, None, grad_in, None, None, None
