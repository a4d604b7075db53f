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
        self._packed, self._packed_key = None, None
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

    # ---- eval mode: each scale is one fused call into libmgar_lidar.so (csrc/voxel_sa.cu) ---------------------------
    @staticmethod
    def _fold(conv, bn):
        w = conv.weight.detach().float().reshape(conv.out_channels, conv.in_channels)
        scale = bn.weight.detach().float() / torch.sqrt(bn.running_var.detach().float() + bn.eps)
        return [w.t().contiguous().reshape(-1), scale, bn.bias.detach().float() - bn.running_mean.detach().float() * scale]

    def _pack(self, device):
        tensors = [t for m in (self.mlps_in, self.mlps_pos, self.mlps_out) for t in list(m.parameters()) + list(m.buffers())]
        key = (str(device),) + tuple((t.data_ptr(), t._version) for t in tensors)
        if self._packed is None or self._packed_key != key:      # re-folded when a weight / running statistic changes
            self._packed = []
            for mi, mp, mo in zip(self.mlps_in, self.mlps_pos, self.mlps_out):
                parts = self._fold(mi[0], mi[1]) + self._fold(mp[0], mp[1]) + self._fold(mo[0], mo[1])
                self._packed.append(((mi[0].in_channels, mi[0].out_channels, mo[0].out_channels),
                                     torch.cat(parts).contiguous().to(device)))
            self._packed_key = key
        return self._packed

    def _fused_ok(self, features):
        from ..... import _lib
        if self.training or torch.is_grad_enabled() or not features.is_cuda or features.dtype != torch.float32:
            return False
        return all(_lib.lib().mgar_voxel_sa_supported(mi[0].in_channels, mi[0].out_channels, mo[0].out_channels, g.nsample)
                   for mi, mo, g in zip(self.mlps_in, self.mlps_out, self.groupers))

    def _forward_fused(self, xyz, new_xyz, new_coords, features, voxel2point_indices):
        from ..... import _lib
        from ....._torch_glue import on_device, ptr, stream_of
        packed = self._pack(features.device)
        N, M = features.shape[0], new_xyz.shape[0]
        out = torch.empty((M, sum(p[0][2] for p in packed)), dtype=torch.float32, device=features.device)
        features, xyz, new_xyz = features.contiguous(), xyz.contiguous(), new_xyz.contiguous()
        col = 0
        for grouper, ((c_in, c1, c2), params) in zip(self.groupers, packed):
            idx, empty = voxel_query_utils.voxel_query(grouper.max_range, grouper.radius, grouper.nsample, xyz, new_xyz,
                                                       new_coords, voxel2point_indices)      # global rows, empty rows zeroed
            em = empty.to(torch.uint8)
            L = _lib.lib()
            nb = int(L.mgar_voxel_sa_workspace_bytes(N, c1))
            ws = torch.empty((max(nb, 16),), dtype=torch.uint8, device=features.device)
            view = out[:, col:col + c2]
            with on_device(features):
                _lib.check(L.mgar_voxel_sa_fwd(N, M, c_in, c1, c2, grouper.nsample, 1 if self.pool_method == 'avg_pool' else 0,
                                               ptr(features), ptr(xyz), ptr(new_xyz), ptr(idx), ptr(em), ptr(params),
                                               view.data_ptr(), out.stride(0), ptr(ws), nb, stream_of(features)), "voxel_sa_fwd")
            col += c2
        return out

    def forward(self, xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt, new_coords, features, voxel2point_indices):
        if self.pool_method not in ('max_pool', 'avg_pool'):
            raise NotImplementedError
        new_coords = new_coords[:, [0, 3, 2, 1]].contiguous()                  # -> [batch, z, y, x]
        if self._fused_ok(features):
            return self._forward_fused(xyz, new_xyz, new_coords, features, voxel2point_indices)
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
