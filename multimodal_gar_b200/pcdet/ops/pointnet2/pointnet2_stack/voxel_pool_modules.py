"""Host-side mirror of /root/reference/pcdet/ops/pointnet2/pointnet2_stack/voxel_pool_modules.py:
`NeighborVoxelSAModuleMSG` with the reference's sub-module names (`groupers`, `mlps_in`, `mlps_pos`, `mlps_out`), so
its checkpoints load unchanged."""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import voxel_query_utils


def _conv_bn(conv, c_in, c_out, bn, relu=False):
    layers = [conv(c_in, c_out, kernel_size=1, bias=False), bn(c_out)]
    return nn.Sequential(*(layers + [nn.ReLU()] if relu else layers))


class NeighborVoxelSAModuleMSG(nn.Module):
    """reference :8-130.  Per scale k: features -> mlps_in[k] (1x1 conv + BN on the N points, BEFORE grouping),
    voxel query + grouping, + mlps_pos[k] of the centred neighbour offsets, ReLU, max/avg over the samples,
    mlps_out[k]; the scales are concatenated.  forward(xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt,
    new_coords (M,4) [batch,x,y,z], features (N,C), voxel2point_indices (B,Z,Y,X)) -> (M, sum_k mlps[k][2])."""

    def __init__(self, *, query_ranges, radii, nsamples, mlps, use_xyz=True, pool_method='max_pool'):
        super().__init__()
        assert len(query_ranges) == len(nsamples) == len(mlps)
        self.groupers, self.mlps_in, self.mlps_pos, self.mlps_out = (nn.ModuleList() for _ in range(4))
        for rng, radius, nsample, spec in zip(query_ranges, radii, nsamples, mlps):
            self.groupers.append(voxel_query_utils.VoxelQueryAndGrouping(rng, radius, nsample))
            self.mlps_in.append(_conv_bn(nn.Conv1d, spec[0], spec[1], nn.BatchNorm1d))
            self.mlps_pos.append(_conv_bn(nn.Conv2d, 3, spec[1], nn.BatchNorm2d))
            self.mlps_out.append(_conv_bn(nn.Conv1d, spec[1], spec[2], nn.BatchNorm1d, relu=True))
        self.relu = nn.ReLU()
        self.pool_method = pool_method
        self.init_weights()

    def init_weights(self):
        for m in self.modules():
            if isinstance(m, (nn.Conv1d, nn.Conv2d)):
                nn.init.kaiming_normal_(m.weight)
                if m.bias is not None:
                    nn.init.constant_(m.bias, 0)
            elif isinstance(m, (nn.BatchNorm1d, nn.BatchNorm2d)):
                nn.init.constant_(m.weight, 1.0)
                nn.init.constant_(m.bias, 0)

    def forward(self, xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt, new_coords, features, voxel2point_indices):
        if self.pool_method not in ('max_pool', 'avg_pool'):
            raise NotImplementedError
        new_coords = new_coords[:, [0, 3, 2, 1]].contiguous()                  # -> [batch, z, y, x]
        scales = []
        for grouper, mlp_in, mlp_pos, mlp_out in zip(self.groupers, self.mlps_in, self.mlps_pos, self.mlps_out):
            f_in = mlp_in(features.t().unsqueeze(0))[0].t().contiguous()        # (N, C1)
            g_feat, g_xyz, empty = grouper(new_coords, xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt, f_in,
                                           voxel2point_indices)               # (M,C1,ns), (M,3,ns), (M)
            gone = empty.view(-1, 1, 1)
            g_feat = g_feat.masked_fill(gone, 0)
            offs = (g_xyz - new_xyz.unsqueeze(-1)).masked_fill(gone, 0)
            pos = mlp_pos(offs.permute(1, 0, 2).unsqueeze(0))                   # (1, C1, M, ns)
            h = self.relu(g_feat.permute(1, 0, 2).unsqueeze(0) + pos)
            h = h.max(dim=3).values if self.pool_method == 'max_pool' else h.mean(dim=3)   # (1, C1, M)
            scales.append(mlp_out(h)[0].t())                                    # (M, C2)
        return torch.cat(scales, dim=1)
