"""Output contract of the LiDAR branch and the input path of the Distance Aware Fusion Module.

Reference: `LiDAR_Backbone.forward` (/root/reference/model/gat_model.py:868-971) returns per-person
features `[B, MAX_NUM_PROPOSAL, 512]` through an `nn.Linear(..., 512)` called `embedding`
(:883-894, :899-912); `GAR_Fusion_Net3.forward` (:1354-1405, :1467-1475) then, in a Python loop over
frames, slices `[:person_num[b]]` (person_num = len(unique(person_id[b])) - 1, :1276-1280), applies
`bn_lidar` and builds De = pairwise_euclidean_distance(bboxes3d[b, :n, :3], zero_diagonal=True).

Here the point-based branch (`lidar_branch.PointPersonEncoder`, the hot path) feeds the same contract and
the per-frame loop is replaced by batched tensors with a validity mask; every frame's `[:n]` slice of the
batched result equals what the reference loop computes for that frame (tests/test_host_logic.py).
"""
import torch
import torch.nn as nn

from . import lidar_branch


def count_persons(person_id):
    """Distinct ids per frame minus one (the padding id), on the device: (B,MNP) -> (B) int64.  One sort of the batch
    instead of the reference's `torch.unique` per frame (64 kernels + 64 host round trips at configs[2])."""
    if person_id.shape[1] == 0:
        return torch.full((person_id.shape[0],), -1, dtype=torch.int64, device=person_id.device)
    s, _ = torch.sort(person_id, dim=1)
    return (s[:, 1:] != s[:, :-1]).sum(dim=1)


def get_num_person(person_id):
    """gat_model.py:1276-1280: persons per frame = distinct ids minus the padding id.  (B,MNP) -> list[int]."""
    return count_persons(person_id).tolist()


def pairwise_euclidean_distance(x, zero_diagonal=True):
    """torchmetrics.functional.pairwise_euclidean_distance restated (the dependency is unpinned and absent,
    SURVEY 8c): sqrt(clamp(|x|^2 + |y|^2 - 2 x.y^T, 0)), diagonal zeroed.  x (..., n, d) -> (..., n, n)."""
    xn = (x * x).sum(dim=-1, keepdim=True)
    d2 = xn + xn.transpose(-1, -2) - 2.0 * (x @ x.transpose(-1, -2))
    d = d2.clamp_min(0).sqrt()
    if zero_diagonal:
        n = x.shape[-2]
        d = d.masked_fill(torch.eye(n, dtype=torch.bool, device=x.device), 0.0)
    return d


class _DafmInputs(dict):
    """The dict `dafm_lidar_inputs` returns; 'person_num' (list[int]) is read back from the device only when somebody
    asks for it - the fused path needs the counts on the device only ('person_count'), and a host read in every step
    stalls the launch queue (the training step lost a third of its time to it)."""

    def __missing__(self, key):
        if key != 'person_num':
            raise KeyError(key)
        self[key] = self['person_count'].tolist()
        return self[key]


def dafm_lidar_inputs(lidar_feature, bboxes3d, person_id):
    """Batched form of the per-frame slicing at gat_model.py:1396-1405 and :1467-1471.

    -> dict(person_count (B) int64 on the device, person_num list[int] (read back on first access), valid (B,MNP) bool,
            feature (B,MNP,D) with padded rows zeroed, De (B,MNP,MNP) with rows/columns of padded persons zeroed).
    `out['feature'][b, :n]` and `out['De'][b, :n, :n]` are the reference's LiDAR_feature_b and De."""
    B, MNP = person_id.shape
    n = count_persons(person_id).to(lidar_feature.device)
    valid = torch.arange(MNP, device=lidar_feature.device).view(1, MNP) < n.view(B, 1)
    feat = lidar_feature * valid.unsqueeze(-1).to(lidar_feature.dtype)
    De = pairwise_euclidean_distance(bboxes3d[..., :3])
    pair = (valid.unsqueeze(1) & valid.unsqueeze(2)).to(De.dtype)
    return _DafmInputs(person_count=n, valid=valid, feature=feat, De=De * pair)


class PointLiDARBackbone(nn.Module):
    """Point-based `LiDAR_Backbone`: data_dict -> shared feature `[B, MAX_NUM_PROPOSAL, 512]`.

    data_dict: 'points' (B,N,3+C) float32, 'gt_bboxes3d' (B,MNP,7) (zero-padded rows allowed: they pool no
    points and get the embedding of a zero vector, exactly like any empty person; the fusion net slices them
    away).  The 128 -> 512 `embedding` runs on tcgen05 (csrc/gemm_tc.cu, 3xTF32) in eval mode without autograd, through
    torch's fp32 GEMM otherwise."""

    def __init__(self, point_feature_channels=1, num_sampled_points=512, npoint=512, radius=0.3, nsample=32,
                 mlp=(64, 128), out_dim=512):
        super().__init__()
        # proposals arrive zero-padded to MAX_NUM_PROPOSAL: the padded slots pool nothing and are skipped by the MLP
        self.encoder = lidar_branch.PointPersonEncoder(point_feature_channels, num_sampled_points, npoint, radius,
                                                       nsample, mlp, skip_empty_persons=True)
        self.embedding = nn.Linear(mlp[-1], out_dim)

    def forward(self, data_dict):
        pts = data_dict['points']
        boxes = data_dict['gt_bboxes3d']
        person = self.encoder(pts[..., 0:3].contiguous(), pts[..., 3:].contiguous(), boxes.contiguous())
        from . import tc_gemm
        if not self.training and tc_gemm.usable(person, self.embedding.out_features, self.embedding.in_features):
            return tc_gemm.linear(person, self.embedding.weight, self.embedding.bias)     # tcgen05, fp32-faithful
        return self.embedding(person)                                   # (B, MNP, out_dim)
