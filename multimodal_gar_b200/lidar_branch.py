"""The LiDAR hot path of MGAR-net composed from the drop-in pcdet.ops surface: per-person point
pooling -> canonical frame -> farthest point sampling -> ball query + grouping + shared MLP ->
per-person feature, i.e. what a point-based `LiDAR_Backbone` (/root/reference/model/gat_model.py:868-971)
hands to the Distance Aware Fusion Module as `[B, NP, C]` (gat_model.py:1359-1360).

The composition follows the reference's own point pooling head (PointRCNNHead.roipool3d_gpu,
/root/reference/pcdet/models/roi_heads/pointrcnn_head.py:85-130) followed by a stack-layout SA
module (pointnet2_stack/pointnet2_modules.py:78-112) with every person as one "batch" of the
stack layout (xyz_batch_cnt = [S] * persons).

Frames are independent: `shard_frames` hands each rank a contiguous block of frames and no
collective is needed in the forward pass.
"""
import os

import torch
import torch.nn as nn

from . import fastpath
from .pcdet.ops.pointnet2.pointnet2_stack import pointnet2_modules, pointnet2_utils, sa_mlp_fused
from .pcdet.ops.roiaware_pool3d import roiaware_pool3d_utils
from .pcdet.ops.roipoint_pool3d import roipoint_pool3d_cuda, roipoint_pool3d_utils
from .pcdet.utils import box_utils, common_utils


def shard_frames(num_frames, world_size, rank):
    """Contiguous block partition of frames over ranks: frame f -> rank floor(f * W / B).
    Returns (first, last_exclusive).  Blocks differ by at most one frame."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad rank %d / world %d" % (rank, world_size))
    lo = (rank * num_frames + world_size - 1) // world_size
    hi = ((rank + 1) * num_frames + world_size - 1) // world_size
    return lo, hi


def canonical_transform(pooled_features, rois, pooled_empty_flag):
    """pointrcnn_head.py:121-129 with the reference's torch ops: (B,M,S,3+C) -> (B*M,S,3+C)."""
    pooled_features = pooled_features.clone()
    pooled_features[:, :, :, 0:3] -= rois[:, :, 0:3].unsqueeze(dim=2)
    pooled_features = pooled_features.view(-1, pooled_features.shape[-2], pooled_features.shape[-1])
    pooled_features[:, :, 0:3] = common_utils.rotate_points_along_z(
        pooled_features[:, :, 0:3], -rois.view(-1, rois.shape[-1])[:, 6])
    pooled_features[pooled_empty_flag.view(-1) > 0] = 0
    return pooled_features


class PointPersonEncoder(nn.Module):
    """points (B,N,3) + point_features (B,N,C) + person boxes (B,M,7) -> person features (B,M,C_out).

    Defaults are BASELINE.json configs[1]: S=512 pooled points, P=512 FPS centroids, ball query
    r=0.3 / nsample=32, shared MLP [C,64,128] (+3 for use_xyz), max over samples, max over centroids.
    """

    def __init__(self, point_feature_channels=1, num_sampled_points=512, npoint=512, radius=0.3, nsample=32,
                 mlp=(64, 128), pool_extra_width=(0.0, 0.0, 0.0), fused_canonical=True, skip_empty_persons=True):
        super().__init__()
        # True: the tensor-core MLP walks only the persons that pooled points (same feature bits; a padded or empty
        # person slot costs nothing; with no empty person the walk measures on par with the plain one)
        self.skip_empty_persons = skip_empty_persons
        # True: ... and of those persons only the distinct centroids (a person that pooled c < P points has c distinct
        # samples; farthest point sampling returns index 0 for the rest, whose features repeat centroid 0's)
        self.skip_repeated_centroids = os.environ.get('MGAR_TC_NO_LIVE_TILES', '0') == '0'
        # True: ... and of those centroids only the distinct samples of each ball (a ball row repeats its hits once it
        # runs out of distinct ones: the copies of the wrapped cloud, then the first hit as padding - same maxima)
        self.skip_repeated_samples = os.environ.get('MGAR_TC_NO_LIVE_UNITS', '0') == '0'
        self.num_sampled_points, self.npoint = num_sampled_points, npoint
        self.pool_extra_width = tuple(pool_extra_width)
        self.fused_canonical = fused_canonical
        self.roipoint_pool3d_layer = roipoint_pool3d_utils.RoIPointPool3d(num_sampled_points, self.pool_extra_width)
        self.sa = pointnet2_modules.StackSAModuleMSG(radii=[radius], nsamples=[nsample],
                                                     mlps=[[point_feature_channels] + list(mlp)], use_xyz=True,
                                                     pool_method='max_pool')
        self.out_channels = mlp[-1]
        # the (B*M*P, C_out) per-centroid SA features are an intermediate of the person feature: stored only on
        # request (return_stages=True or this flag); the max over a person's centroids happens in the MLP kernel
        self.keep_sa_features = os.environ.get('MGAR_KEEP_SA', '0') != '0'

    @torch.no_grad()
    def pool(self, points, point_features, boxes):
        """-> canonical pooled (B*M,S,3+C), empty flag (B,M) int32."""
        B, M = boxes.shape[0], boxes.shape[1]
        C, S = point_features.shape[2], self.num_sampled_points
        if self.fused_canonical:
            # one kernel: in-box test + ordered compaction + gather + canonical frame
            big = box_utils.enlarge_box3d(boxes.reshape(-1, 7), self.pool_extra_width).view(B, M, 7).contiguous()
            pooled = torch.empty((B, M, S, 3 + C), dtype=torch.float32, device=points.device)
            empty = torch.empty((B, M), dtype=torch.int32, device=points.device)
            roipoint_pool3d_cuda.forward(points.contiguous(), big, point_features.contiguous(), pooled, empty,
                                         canonical=True)
            return pooled.view(B * M, S, 3 + C), empty
        pooled, empty = self.roipoint_pool3d_layer(points, point_features, boxes)
        return canonical_transform(pooled, boxes, empty), empty

    def forward(self, points, point_features, boxes, return_stages=False, mark=None):
        """`mark(name)` (optional) is called after each stage has been enqueued - bench.py records a
        CUDA event there to time the stages inside the step."""
        user_mark = mark or (lambda name: None)
        nvtx = torch.cuda.nvtx if points.is_cuda else None

        def mark(name):
            # one NVTX range per stage (the profiler hooks of the path) - a stage ends where `mark` is called
            if nvtx is not None:
                nvtx.range_pop()
                nvtx.range_push("mgar:after_" + name)
            user_mark(name)
        if nvtx is not None:
            nvtx.range_push("mgar:lidar_branch")
            nvtx.range_push("mgar:begin")
        try:
            return self._forward(points, point_features, boxes, return_stages, mark)
        finally:
            if nvtx is not None:
                nvtx.range_pop()
                nvtx.range_pop()

    def _forward(self, points, point_features, boxes, return_stages, mark):
        B, M = boxes.shape[0], boxes.shape[1]
        S, P = self.num_sampled_points, self.npoint
        if self.fused_canonical and os.environ.get("MGAR_POOL", "mask") == "mask":
            # one pass over the points for all boxes (first-hit index + in-box bit masks), then per-box gather
            C = point_features.shape[2]
            same = all(w == 0 for w in self.pool_extra_width)
            big = boxes if same else box_utils.enlarge_box3d(boxes.reshape(-1, 7), self.pool_extra_width).view(B, M, 7).contiguous()
            box_idx, mask = fastpath.box_masks(points.contiguous(), boxes.contiguous(), None if same else big)
            mark('points_in_boxes')
            pooled, empty, pooled_cnt = fastpath.roipoint_pool_from_masks(points.contiguous(), point_features.contiguous(),
                                                                          big.contiguous(), mask, S, want_cnt=True)
            pooled = pooled.view(B * M, S, 3 + C)
            mark('roipoint_pool')
        else:
            box_idx = roiaware_pool3d_utils.points_in_boxes_gpu(points, boxes)      # (B,N) first box or -1
            mark('points_in_boxes')
            pooled, empty = self.pool(points, point_features, boxes)                # (B*M,S,3+C) canonical frame
            mark('roipoint_pool')
            pooled_cnt = None
        g = self.sa.groupers[0]
        ball_idx = live_cnt = None
        packed = None if (self.training or torch.is_grad_enabled()) else self.sa._pack(pooled.device)[0]
        C = pooled.shape[2] - 3
        want_out = return_stages or self.keep_sa_features
        # nsample 16 / 64 run on the tensor cores too when only the person max is wanted (blocks of 32 positions)
        if packed is not None and (sa_mlp_fused.tc_supported(packed[0], g.nsample, C) or
                                   (not want_out and (P * g.nsample) % 32 == 0 and
                                    sa_mlp_fused.tc_group_max_supported(packed[0], g.nsample, C))):
            # fused path: every kernel reads the pooled (.., 3+C) rows in place; 5 launches per step
            flat = pooled.view(B * M * S, 3 + C)
            # distinct-centroid walk: pooled clouds repeat their points, the sampling kernel reports how many distinct
            # samples each person has and the MLP computes only those (the others repeat centroid 0: same max)
            walk = self.skip_empty_persons and self.skip_repeated_centroids and not want_out
            fused = fastpath.fps_ball_query(pooled[:, :, 0:3], P, g.radius, g.nsample, all_rows=not walk,
                                            cloud_weight=None if pooled_cnt is None else pooled_cnt.view(-1))
            if fused is not None:                  # sampling and ball query share their distances: one kernel
                fps_idx, new_xyz, ball_idx, live_cnt = fused
                new_xyz = new_xyz.view(B * M * P, 3)
                mark('fps_ball_query')
            else:
                fps_idx, new_xyz = fastpath.fps_gather(pooled[:, :, 0:3], P)
                new_xyz = new_xyz.view(B * M * P, 3)
                mark('fps')
                ball_idx = fastpath.ball_query_uniform(g.radius, g.nsample, flat[:, 0:3], new_xyz, S, P)
                mark('ball_query')
            person = torch.empty((B * M, self.out_channels), dtype=torch.float32, device=pooled.device)
            sa_feat = sa_mlp_fused.sa_mlp_forward(packed[0], packed[1], g.nsample, flat[:, 0:3], None, new_xyz, None,
                                                  flat[:, 3:] if C else None, ball_idx, uniform=(S, P),
                                                  params_host=packed[2], impl='tc', group_max=person, group_size=P,
                                                  group_zero=empty.view(-1), want_out=want_out,
                                                  skip_zero_groups=self.skip_empty_persons,
                                                  group_live_cnt=live_cnt if walk else None,
                                                  group_point_cnt=(pooled_cnt.view(-1) if (walk and self.skip_repeated_samples and
                                                                                          pooled_cnt is not None) else None))
            mark('sa_mlp')
            person = person.view(B, M, -1)
        else:
            xyz = pooled[:, :, 0:3].contiguous()                                    # (B*M,S,3)
            feats = pooled[:, :, 3:].contiguous().view(B * M * S, -1)
            fps_idx = pointnet2_utils.farthest_point_sample(xyz, P)                 # (B*M,P) int32
            mark('fps')
            new_xyz = torch.gather(xyz, 1, fps_idx.long().unsqueeze(-1).expand(-1, -1, 3)).view(B * M * P, 3)
            cnt = torch.full((B * M,), S, dtype=torch.int32, device=xyz.device)
            new_cnt = torch.full((B * M,), P, dtype=torch.int32, device=xyz.device)
            self.sa.stage_mark = mark
            _, sa_feat = self.sa(xyz.view(B * M * S, 3), cnt, new_xyz, new_cnt, feats)  # (B*M*P, C_out)
            self.sa.stage_mark = None
            mark('sa_mlp')
            person = sa_feat.view(B * M, P, -1).max(dim=1).values
            person = person * (empty.view(-1, 1) == 0).to(person.dtype)             # empty persons -> zeros
            person = person.view(B, M, -1)
            mark('person_reduce')
        if return_stages:
            # ball_idx: raw kernel rows (cloud-local indices, -1 in slot 0 of an empty ball), fused path only
            return dict(box_idx=box_idx, pooled=pooled, empty=empty, fps_idx=fps_idx, new_xyz=new_xyz, ball_idx=ball_idx,
                        live_cnt=live_cnt,
                        sa_features=sa_feat, person_features=person)
        return person
