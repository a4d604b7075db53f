"""Host-side mirror of /root/reference/pcdet/ops/pointnet2/pointnet2_batch/pointnet2_modules.py:
`PointnetSAModuleMSG`, `PointnetSAModule`, `PointnetFPModule` with the reference's sub-module names (`groupers`,
`mlps`, `mlp`), so PointNet2MSG / PointRCNN checkpoints load unchanged.  The shared MLPs are library convolutions."""
import torch
import torch.nn as nn

from . import pointnet2_utils


def _shared_mlp(spec):
    layers = []
    for c_in, c_out in zip(spec, spec[1:]):
        layers += [nn.Conv2d(c_in, c_out, kernel_size=1, bias=False), nn.BatchNorm2d(c_out), nn.ReLU()]
    return nn.Sequential(*layers)


class _PointnetSAModuleBase(nn.Module):
    def __init__(self):
        super().__init__()
        self.npoint = None
        self.groupers = None
        self.mlps = None
        self.pool_method = 'max_pool'

    def forward(self, xyz, features=None, new_xyz=None):
        """reference :19-55.  xyz (B,N,3), features (B,C,N) -> new_xyz (B,npoint,3), (B, sum_k mlps[k][-1], npoint)."""
        if self.pool_method not in ('max_pool', 'avg_pool'):
            raise NotImplementedError
        if new_xyz is None and self.npoint is not None:
            picked = pointnet2_utils.farthest_point_sample(xyz, self.npoint)
            new_xyz = pointnet2_utils.gather_operation(xyz.transpose(1, 2).contiguous(), picked).transpose(1, 2).contiguous()
        scales = []
        for grouper, mlp in zip(self.groupers, self.mlps):
            h = mlp(grouper(xyz, new_xyz, features))                                     # (B, C_out, npoint, nsample)
            scales.append(h.max(dim=3).values if self.pool_method == 'max_pool' else h.mean(dim=3))
        return new_xyz, torch.cat(scales, dim=1)


class PointnetSAModuleMSG(_PointnetSAModuleBase):
    """reference :58-99 (multi-scale grouping; npoint None = one group over the whole cloud).  As in the reference,
    use_xyz adds 3 to mlps[k][0] IN PLACE."""

    def __init__(self, *, npoint, radii, nsamples, mlps, bn=True, use_xyz=True, pool_method='max_pool'):
        super().__init__()
        assert len(radii) == len(nsamples) == len(mlps)
        self.npoint = npoint
        self.groupers, self.mlps = nn.ModuleList(), nn.ModuleList()
        for radius, nsample, spec in zip(radii, nsamples, mlps):
            self.groupers.append(pointnet2_utils.QueryAndGroup(radius, nsample, use_xyz=use_xyz) if npoint is not None
                                 else pointnet2_utils.GroupAll(use_xyz))
            if use_xyz:
                spec[0] += 3
            self.mlps.append(_shared_mlp(spec))
        self.pool_method = pool_method


class PointnetSAModule(PointnetSAModuleMSG):
    """reference :102-119: the single-scale form."""

    def __init__(self, *, mlp, npoint=None, radius=None, nsample=None, bn=True, use_xyz=True, pool_method='max_pool'):
        super().__init__(mlps=[mlp], npoint=npoint, radii=[radius], nsamples=[nsample], bn=bn, use_xyz=use_xyz,
                         pool_method=pool_method)


class PointnetFPModule(nn.Module):
    """reference :122-170: inverse-distance interpolation of known_feats (B,C2,m) onto the unknown points (B,n,3),
    concatenated with unknow_feats (B,C1,n), then the shared MLP -> (B, mlp[-1], n)."""

    def __init__(self, *, mlp, bn=True):
        super().__init__()
        self.mlp = _shared_mlp(mlp)

    def forward(self, unknown, known, unknow_feats, known_feats):
        if known is not None:
            dist, idx = pointnet2_utils.three_nn(unknown, known)
            recip = 1.0 / (dist + 1e-8)
            weight = recip / torch.sum(recip, dim=2, keepdim=True)
            feats = pointnet2_utils.three_interpolate(known_feats, idx, weight)
        else:
            feats = known_feats.expand(*known_feats.size()[0:2], unknown.size(1))
        if unknow_feats is not None:
            feats = torch.cat([feats, unknow_feats], dim=1)
        return self.mlp(feats.unsqueeze(-1)).squeeze(-1)
