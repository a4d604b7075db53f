"""Host-side mirror of /root/reference/pcdet/ops/roipoint_pool3d/roipoint_pool3d_utils.py."""
import torch
import torch.nn as nn
from torch.autograd import Function

from ...utils import box_utils
from . import roipoint_pool3d_cuda


class RoIPointPool3dFunction(Function):
    @staticmethod
    def forward(ctx, points, point_features, boxes3d, pool_extra_width, num_sampled_points=512):
        """reference :33-59.  points (B,N,3), point_features (B,N,C), boxes3d (B,M,7) ->
        pooled_features (B,M,S,3+C), pooled_empty_flag (B,M) int32."""
        assert len(points.shape) == 3 and points.shape[2] == 3
        B, M, C = points.shape[0], boxes3d.shape[1], point_features.shape[2]
        pooled_boxes3d = box_utils.enlarge_box3d(boxes3d.reshape(-1, 7), pool_extra_width).view(B, -1, 7)
        pooled_features = point_features.new_zeros((B, M, num_sampled_points, 3 + C))
        pooled_empty_flag = torch.zeros((B, M), dtype=torch.int32, device=points.device)
        roipoint_pool3d_cuda.forward(points.contiguous(), pooled_boxes3d.contiguous(),
                                     point_features.contiguous(), pooled_features, pooled_empty_flag)
        ctx.mark_non_differentiable(pooled_empty_flag)
        return pooled_features, pooled_empty_flag

    @staticmethod
    def backward(ctx, *grad_out):
        raise NotImplementedError  # as the reference (:61-63)


class RoIPointPool3d(nn.Module):
    """reference :9-28."""

    def __init__(self, num_sampled_points=512, pool_extra_width=1.0):
        super().__init__()
        self.num_sampled_points = num_sampled_points
        self.pool_extra_width = pool_extra_width

    def forward(self, points, point_features, boxes3d):
        return RoIPointPool3dFunction.apply(points, point_features, boxes3d, self.pool_extra_width,
                                            self.num_sampled_points)
