"""Host-side mirror of StackSAModuleMSG from
/root/reference/pcdet/ops/pointnet2/pointnet2_stack/pointnet2_modules.py:30-112 (same constructor,
parameter names `mlps.{k}.{0,1,3,4}.*`, init and forward signature).

eval(): each scale runs as ball query + ONE fused kernel (gather, MLP, BatchNorm folded, ReLU,
pool).  train(): ball query + ONE autograd node per scale (sa_mlp_fused.SAMlpTrainFunction: batch-statistics
BatchNorm forward and the whole backward in libmgar_lidar.so, running statistics updated as nn.BatchNorm2d does).
eval() with autograd enabled, or MGAR_SA_TRAIN=torch: the reference's own composition (QueryAndGroup ->
Conv2d/BN/ReLU -> pool) on the new grouping kernels.
"""
import os
from typing import List

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import pointnet2_utils, sa_mlp_fused


def build_local_aggregation_module(input_channels, config):
    """reference :10-27 (StackSAModuleMSG only; VectorPool is outside this build)."""
    name = config.get('NAME', 'StackSAModuleMSG')
    if name != 'StackSAModuleMSG':
        raise NotImplementedError(name)
    mlps = config.MLPS
    for k in range(len(mlps)):
        mlps[k] = [input_channels] + mlps[k]
    layer = StackSAModuleMSG(radii=config.POOL_RADIUS, nsamples=config.NSAMPLE, mlps=mlps, use_xyz=True,
                             pool_method='max_pool')
    return layer, sum(x[-1] for x in mlps)


class StackSAModuleMSG(nn.Module):
    def __init__(self, *, radii: List[float], nsamples: List[int], mlps: List[List[int]],
                 use_xyz: bool = True, pool_method='max_pool'):
        super().__init__()
        assert len(radii) == len(nsamples) == len(mlps)
        self.groupers = nn.ModuleList()
        self.mlps = nn.ModuleList()
        for radius, nsample, spec in zip(radii, nsamples, mlps):
            self.groupers.append(pointnet2_utils.QueryAndGroup(radius, nsample, use_xyz=use_xyz))
            if use_xyz:
                spec[0] += 3  # in place, as the reference does (:53-54)
            layers = []
            for k in range(len(spec) - 1):
                layers += [nn.Conv2d(spec[k], spec[k + 1], kernel_size=1, bias=False),
                           nn.BatchNorm2d(spec[k + 1]), nn.ReLU()]
            self.mlps.append(nn.Sequential(*layers))
        self.pool_method = pool_method
        self.use_xyz = use_xyz
        self._packed = None
        self._packed_key = None
        self.stage_mark = None   # optional callable(name): profiling hook between ball query and the MLP
        self.init_weights()

    def init_weights(self):
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight)
                if m.bias is not None:
                    nn.init.constant_(m.bias, 0)
            if isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1.0)
                nn.init.constant_(m.bias, 0)

    def train(self, mode=True):
        # parameters may change in train(): re-fold on the next eval forward.  A repeated eval() keeps the folded tensors:
        # a CUDA graph captured over the eval path (multimodal_gar_b200/graphs.py) holds their device addresses, and the
        # version key below catches every real change of a parameter or buffer anyway.
        if mode or self.training:
            self._packed = None
        return super().train(mode)

    def _pack_key(self, device):
        # every tensor the folded parameters are made of: storage and in-place version, so that load_state_dict(), an EMA
        # swap or any in-place edit of a weight / running statistic re-folds (a stale cache would silently compute with
        # the old numbers)
        key = [str(device)]
        for t in list(self.mlps.parameters()) + list(self.mlps.buffers()):
            key.append((t.data_ptr(), t._version))
        return tuple(key)

    def _pack(self, device):
        key = self._pack_key(device)
        if self._packed is None or self._packed_key != key:
            self._packed = []
            for m in self.mlps:
                channels, params = sa_mlp_fused.pack_shared_mlp(m)
                host = params.cpu().contiguous()
                self._packed.append((channels, host.to(device), host))
            self._packed_key = key
        return self._packed

    def forward(self, xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt, features=None, empty_voxel_set_zeros=True):
        """reference :78-112 -> (new_xyz, new_features (M, sum_k mlps[k][-1]))."""
        fused = (not self.training) and self.use_xyz and self.pool_method in ('max_pool', 'avg_pool') \
            and not torch.is_grad_enabled()
        if fused:
            packed = self._pack(xyz.device)
            widths = [p[0][-1] for p in packed]
            out = torch.empty((new_xyz.shape[0], sum(widths)), dtype=torch.float32, device=xyz.device)
            col = 0
            for k, (channels, params, params_host) in enumerate(packed):
                g = self.groupers[k]
                idx, empty = pointnet2_utils.ball_query(g.radius, g.nsample, xyz, xyz_batch_cnt, new_xyz,
                                                        new_xyz_batch_cnt)
                if self.stage_mark is not None:
                    self.stage_mark('ball_query')
                sa_mlp_fused.sa_mlp_forward(channels, params, g.nsample, xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt,
                                            features, idx, empty_mask=empty, out=out[:, col:col + widths[k]],
                                            pool_avg=self.pool_method == 'avg_pool', params_host=params_host)
                col += widths[k]
            return new_xyz, out

        native_train = self.training and self.use_xyz and self.pool_method in ('max_pool', 'avg_pool') \
            and os.environ.get("MGAR_SA_TRAIN", "native") != "torch" and new_xyz.shape[0] > 0
        if native_train:
            # train(): BatchNorm batch statistics + backward through libmgar_lidar (one autograd node per scale)
            outs = []
            for k, g in enumerate(self.groupers):
                idx, empty = pointnet2_utils.ball_query(g.radius, g.nsample, xyz, xyz_batch_cnt, new_xyz,
                                                        new_xyz_batch_cnt)
                outs.append(sa_mlp_fused.sa_mlp_train_forward(self.mlps[k], g.nsample, xyz.contiguous(), xyz_batch_cnt,
                                                              new_xyz.contiguous(), new_xyz_batch_cnt, features, idx,
                                                              empty_mask=empty, pool_avg=self.pool_method == 'avg_pool'))
            return new_xyz, torch.cat(outs, dim=1)

        new_features_list = []
        for k in range(len(self.groupers)):
            new_features, _ = self.groupers[k](xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt, features)
            new_features = new_features.permute(1, 0, 2).unsqueeze(dim=0)      # (1, C, M, nsample)
            new_features = self.mlps[k](new_features)
            if self.pool_method == 'max_pool':
                new_features = F.max_pool2d(new_features, kernel_size=[1, new_features.size(3)]).squeeze(dim=-1)
            elif self.pool_method == 'avg_pool':
                new_features = F.avg_pool2d(new_features, kernel_size=[1, new_features.size(3)]).squeeze(dim=-1)
            else:
                raise NotImplementedError
            new_features_list.append(new_features.squeeze(dim=0).permute(1, 0))  # (M, C)
        return new_xyz, torch.cat(new_features_list, dim=1)


class StackPointnetFPModule(nn.Module):
    """reference :115-157 (same constructor, sub-module names `mlp.{0,1,2,...}` and forward signature): feature
    propagation on ragged batches - three nearest known points (libmgar_lidar three_nn), inverse-distance weights,
    interpolation (libmgar_lidar, autograd through ThreeInterpolate), shared 1x1 MLP (library convolutions)."""

    def __init__(self, *, mlp: List[int]):
        super().__init__()
        shared_mlps = []
        for k in range(len(mlp) - 1):
            shared_mlps.extend([
                nn.Conv2d(mlp[k], mlp[k + 1], kernel_size=1, bias=False),
                nn.BatchNorm2d(mlp[k + 1]),
                nn.ReLU()
            ])
        self.mlp = nn.Sequential(*shared_mlps)

    def forward(self, unknown, unknown_batch_cnt, known, known_batch_cnt, unknown_feats=None, known_feats=None):
        """unknown (N1+N2..,3), known (M1+M2..,3), unknown_feats (N,C1) or None, known_feats (M,C2) -> (N, C_out)."""
        dist, idx = pointnet2_utils.three_nn(unknown, unknown_batch_cnt, known, known_batch_cnt)
        dist_recip = 1.0 / (dist + 1e-8)
        norm = torch.sum(dist_recip, dim=-1, keepdim=True)
        weight = dist_recip / norm
        interpolated_feats = pointnet2_utils.three_interpolate(known_feats, idx, weight)
        if unknown_feats is not None:
            new_features = torch.cat([interpolated_feats, unknown_feats], dim=1)   # (N, C2 + C1)
        else:
            new_features = interpolated_feats
        new_features = new_features.permute(1, 0)[None, :, :, None]                # (1, C, N, 1)
        new_features = self.mlp(new_features)
        return new_features.squeeze(dim=0).squeeze(dim=-1).permute(1, 0)           # (N, C_out)
