"""Host-side mirror of /root/reference/pcdet/ops/pointnet2/pointnet2_stack/voxel_query_utils.py:
`voxel_query`, `VoxelQueryAndGrouping` - same names, arguments and results, on libmgar_lidar.so."""
import torch
import torch.nn as nn
from torch.autograd import Function

from . import pointnet2_stack_cuda as pointnet2
from . import pointnet2_utils


class VoxelQuery(Function):
    """reference :10-48.  (max_range (z,y,x), radius, nsample, xyz (N,3), new_xyz (M,3), new_coords (M,4) int32
    [batch,z,y,x], point_indices (B,Z,Y,X) int32) -> idx (M,nsample) int32 GLOBAL point indices (rows without a
    neighbour zeroed), empty_ball_mask (M) bool."""

    @staticmethod
    def forward(ctx, max_range, radius, nsample, xyz, new_xyz, new_coords, point_indices):
        for t in (new_xyz, xyz, new_coords, point_indices):
            assert t.is_contiguous()
        M = new_coords.shape[0]
        _, Z, Y, X = point_indices.shape
        idx = torch.zeros((M, nsample), dtype=torch.int32, device=new_xyz.device)
        z_range, y_range, x_range = max_range
        pointnet2.voxel_query_wrapper(M, Z, Y, X, nsample, radius, z_range, y_range, x_range, new_xyz, xyz, new_coords,
                                      point_indices, idx)
        empty_ball_mask = idx[:, 0] == -1
        idx[empty_ball_mask] = 0
        ctx.mark_non_differentiable(idx, empty_ball_mask)
        return idx, empty_ball_mask

    @staticmethod
    def backward(ctx, a=None, b=None):
        return None, None, None, None, None, None, None


voxel_query = VoxelQuery.apply


class VoxelQueryAndGrouping(nn.Module):
    """reference :53-100.  forward(new_coords (M,4) [batch,z,y,x], xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt,
    features (N,C), voxel2point_indices (B,Z,Y,X)) -> grouped_features (M,C,nsample), grouped_xyz (M,3,nsample),
    empty_ball_mask (M).  As in the reference every frame must hold the same number of queries (the
    `view(batch_size, -1, nsample)` at :85)."""

    def __init__(self, max_range, radius, nsample):
        super().__init__()
        self.max_range, self.radius, self.nsample = max_range, radius, nsample

    def forward(self, new_coords, xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt, features, voxel2point_indices):
        assert xyz.shape[0] == xyz_batch_cnt.sum(), 'xyz: %s, xyz_batch_cnt: %s' % (xyz.shape, xyz_batch_cnt)
        assert new_coords.shape[0] == new_xyz_batch_cnt.sum(), \
            'new_coords: %s, new_xyz_batch_cnt: %s' % (new_coords.shape, new_xyz_batch_cnt)
        batch_size = xyz_batch_cnt.shape[0]
        idx, empty_ball_mask = voxel_query(self.max_range, self.radius, self.nsample, xyz, new_xyz, new_coords,
                                           voxel2point_indices)
        # global -> frame-local indices (the grouping kernel adds the frame offset back), one subtraction for all frames
        start = torch.cumsum(xyz_batch_cnt, 0, dtype=torch.int32) - xyz_batch_cnt.to(torch.int32)
        idx = (idx.view(batch_size, -1, self.nsample) - start.view(-1, 1, 1)).view(-1, self.nsample)
        idx[empty_ball_mask] = 0
        grouped_xyz = pointnet2_utils.grouping_operation(xyz, xyz_batch_cnt, idx, new_xyz_batch_cnt)
        grouped_features = pointnet2_utils.grouping_operation(features, xyz_batch_cnt, idx, new_xyz_batch_cnt)
        return grouped_features, grouped_xyz, empty_ball_mask
