"""Stand-in for the reference's `roipoint_pool3d_cuda` pybind module
(/root/reference/pcdet/ops/roipoint_pool3d/src/roipoint_pool3d.cpp:23-59), backed by libmgar_lidar.so."""
import torch

from .... import _lib
from ...._torch_glue import need_cuda, need_dtype, on_device, ptr, stream_of


def forward(xyz, boxes3d, pts_feature, pooled_features, pooled_empty_flag, canonical=False, pts_idx=None):
    """xyz (B,N,3), boxes3d (B,M,7) already enlarged, pts_feature (B,N,C) ->
    pooled_features (B,M,S,3+C), pooled_empty_flag (B,M) i32; S and C are read from the shapes
    (roipoint_pool3d.cpp:33-39).  `canonical` / `pts_idx` are extensions (default: reference behaviour)."""
    need_cuda(xyz=xyz, boxes3d=boxes3d, pts_feature=pts_feature, pooled_features=pooled_features,
              pooled_empty_flag=pooled_empty_flag, pts_idx=pts_idx)
    need_dtype(torch.float32, xyz=xyz, boxes3d=boxes3d, pts_feature=pts_feature, pooled_features=pooled_features)
    need_dtype(torch.int32, pooled_empty_flag=pooled_empty_flag, pts_idx=pts_idx)
    B, N, M = xyz.size(0), xyz.size(1), boxes3d.size(1)
    C, S = pts_feature.size(2), pooled_features.size(2)
    if pooled_features.size(3) != 3 + C:
        raise _lib.MgarError("pooled_features last dim must be 3 + C")
    with on_device(xyz):
        _lib.check(_lib.lib().mgar_roipoint_pool(ptr(xyz), ptr(boxes3d), ptr(pts_feature), ptr(pooled_features),
                                                 ptr(pooled_empty_flag), ptr(pts_idx), B, N, M, S, C,
                                                 1 if canonical else 0, stream_of(xyz)), "roipoint_pool3d forward")
    return 1
