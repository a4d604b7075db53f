"""configs[3]'s share of the hot path: a forward + backward + optimizer step of the LiDAR branch and the Distance Aware
Fusion Module under data parallelism (north_star config #4).

Reference: the training step is /root/reference/train_func.py:98-111,262-269 (model(batch) -> multi-head loss ->
`optimizer.zero_grad(); loss.backward(); optimizer.step()` with Adam, :552) around `GAR_Fusion_ALL` replicated by
`nn.DataParallel` (:512): one process, a replica per GPU, the batch split by frame, gradients summed on GPU 0, BatchNorm
batch statistics per replica (DataParallel does not synchronise them), parameters re-broadcast every step.

Here: one PROCESS per GPU and `DistributedDataParallel` - the gradients are all-reduced with NCCL over NVLink, bucketed
and overlapped with the backward pass; frames are sharded by `lidar_branch.shard_frames`; BatchNorm batch statistics stay
per rank (the reference's semantics); there is no collective in the forward pass.

What is inside the step (and what is not): `PointLiDARBackbone` in train() - the SA shared MLP runs through
`mgar_sa_mlp_fwd_train` / `mgar_sa_mlp_bwd` (BatchNorm batch statistics over all positions, csrc/sa_mlp_train*.cu) -
then the 128 -> 512 embedding, `DistanceAwareFusion` (two `FusionAttention_mat` modules, gat_model.py:427-511,1467-1475)
against the RGB stream, and a linear multi-head classifier with the widths of the reference's person heads (pose 4+4+4,
interaction 2+4+7+5, gat_model.py:1372-1378) trained with a BCE-with-logits loss.  The RGB branch (Inception / RoIAlign)
is outside SURVEY section 8: its output `[B, NP, 512]` arrives as an input tensor (synthetic in bench.py and the tests).
"""
import torch
import torch.nn as nn

from . import lidar_backbone, lidar_branch
from .dafm import DistanceAwareFusion

HEAD_WIDTHS = (4, 4, 4, 2, 4, 7, 5)          # pose_1..3, intrctn_1..4 (gat_model.py:1372-1378)


class LidarFusionNet(nn.Module):
    """batch dict -> scalar loss.  Keys: 'points' (B,N,3+C), 'gt_bboxes3d' (B,MNP,7), 'bboxes' (B,MNP,4),
    'rgb_feature' (B,MNP,D), 'person_id' (B,MNP) int (padding id repeated, as dataloader.py:245-247 pads),
    'labels' (B,MNP,sum(HEAD_WIDTHS)) multi-hot float."""

    def __init__(self, dim=512, sigma=10, backbone=None, **encoder_kw):
        super().__init__()
        # backbone: any module data_dict -> (B, MNP, dim); default = the point-based LiDAR branch (CUDA only)
        self.lidar_backbone = backbone if backbone is not None else lidar_backbone.PointLiDARBackbone(out_dim=dim, **encoder_kw)
        self.fusion = DistanceAwareFusion(dim, sigma)
        self.heads = nn.Linear(dim, sum(HEAD_WIDTHS))

    def features(self, batch):
        lidar = self.lidar_backbone(batch)                                            # (B, MNP, D)
        inp = lidar_backbone.dafm_lidar_inputs(lidar, batch['gt_bboxes3d'], batch['person_id'])
        fused = self.fusion(batch['rgb_feature'], inp['feature'], batch['bboxes'], batch['gt_bboxes3d'], inp['valid'])
        return fused, inp['valid']

    def forward(self, batch):
        fused, valid = self.features(batch)
        logits = self.heads(fused)
        per = nn.functional.binary_cross_entropy_with_logits(logits, batch['labels'], reduction='none').mean(dim=-1)
        w = valid.to(per.dtype)
        return (per * w).sum() / w.sum().clamp_min(1.0)                              # mean over the real persons


def shard_batch(batch, world_size, rank):
    """Frames rank `rank` of `world_size` owns (contiguous blocks)."""
    lo, hi = lidar_branch.shard_frames(next(iter(batch.values())).shape[0], world_size, rank)
    return {k: v[lo:hi].contiguous() for k, v in batch.items()}


def synthetic_batch(seed, B, N=65536, M=80, MNP=100, dim=512, device='cpu'):
    """A JRDB-Act-shaped training batch: the configs[2]/[3] LiDAR frames (M persons zero-padded to MNP proposals) + a
    random RGB feature stream, 2-D boxes, person ids and multi-hot labels."""
    import numpy as np

    from . import synthetic
    pts, boxes = synthetic.make_batch(seed, B, N=N, M=M)
    g = torch.Generator().manual_seed(seed)
    b3 = torch.zeros((B, MNP, 7))
    b3[:, :M] = torch.from_numpy(boxes)
    xy = torch.rand((B, MNP, 2), generator=g) * 600
    wh = torch.rand((B, MNP, 2), generator=g) * 100 + 20
    pid = torch.full((B, MNP), -1, dtype=torch.int64)
    pid[:, :M] = torch.arange(M).view(1, M)
    labels = (torch.rand((B, MNP, sum(HEAD_WIDTHS)), generator=g) < 0.25).float()
    batch = dict(points=torch.from_numpy(np.ascontiguousarray(pts)), gt_bboxes3d=b3,
                 bboxes=torch.cat([xy, xy + wh], dim=-1), rgb_feature=torch.randn((B, MNP, dim), generator=g),
                 person_id=pid, labels=labels)
    return {k: v.to(device) for k, v in batch.items()}


def make_ddp(model, device=None, **kw):
    """`DistributedDataParallel` with the reference's semantics where they matter: buffers (BatchNorm running
    statistics) are NOT broadcast every step (each rank keeps the statistics of its own frames, as each DataParallel
    replica normalises with its own batch statistics); gradients are averaged over ranks in 25 MB buckets that NCCL
    reduces while the backward pass is still running."""
    from torch.nn.parallel import DistributedDataParallel as DDP
    ids = None if device is None or torch.device(device).type != 'cuda' else [torch.device(device).index or 0]
    return DDP(model, device_ids=ids, broadcast_buffers=False, gradient_as_bucket_view=True, **kw)


def grad_bytes(model):
    """Bytes one gradient all-reduce moves per rank (fp32 parameters that require grad)."""
    return sum(p.numel() * p.element_size() for p in model.parameters() if p.requires_grad)


def train_step(model, optimizer, batch):
    """train_func.py:262-269: zero_grad -> forward -> backward (DDP all-reduces inside) -> step.  Returns the loss."""
    optimizer.zero_grad(set_to_none=True)
    loss = model(batch)
    loss.backward()
    optimizer.step()
    return loss.detach()


class GraphedTrainStep:
    """`train_step` for ONE fixed batch shape captured into a CUDA graph: zero_grad + forward + backward + optimizer step
    replayed with one `cudaGraphLaunch`.  The eager step is launch-bound (5.6 ms of kernels in a 6.4 ms step: ~400 small
    launches of the fusion module's forward / backward around the library's few large ones); nothing in it reads the device
    back, so it is capturable as it is.  The optimizer must be capturable (`torch.optim.Adam(..., capturable=True)`); under
    DistributedDataParallel follow torch's rules for graphed DDP (construct the wrapper in a side stream, >= 11 warm-up steps).

        step = GraphedTrainStep(model, optimizer, batch)      # warm-up steps (they DO train) + capture
        loss = step(next_batch)                               # static tensor, overwritten by the next replay"""

    def __init__(self, model, optimizer, batch, warmup=3):
        dev = next(iter(batch.values())).device
        self.model, self.optimizer = model, optimizer
        self.static_batch = {k: v.clone() for k, v in batch.items()}
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                train_step(model, optimizer, self.static_batch)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        optimizer.zero_grad(set_to_none=True)
        with torch.cuda.graph(self.graph):
            self.static_loss = model(self.static_batch)
            self.static_loss.backward()
            optimizer.step()

    def __call__(self, batch):
        for k, dst in self.static_batch.items():
            src = batch[k]
            if dst.data_ptr() != src.data_ptr():
                dst.copy_(src, non_blocking=True)
        self.graph.replay()
        return self.static_loss.detach()
