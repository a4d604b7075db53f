"""configs[3] under data parallelism (multimodal_gar_b200.train_step; reference step train_func.py:262-269 under
nn.DataParallel :512).  CPU part (gloo, world_size 2): the DistributedDataParallel plumbing - frames sharded by
shard_frames, gradients after the all-reduce identical on every rank and equal to the mean of the per-shard
gradients of a single process (BatchNorm-style per-rank statistics are the reference's DataParallel semantics, so the
single-process twin evaluates each shard on its own), optimizer steps keep the replicas bit-identical.  The LiDAR
backbone is CUDA-only: here it is stood in by a differentiable stub with the same output contract; the -m gpu test
below runs the real one (train-mode SA shared MLP through mgar_sa_mlp_fwd_train / mgar_sa_mlp_bwd)."""
import os
import socket
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_WORKER = r"""
import copy, os, sys
import torch
import torch.distributed as dist
import torch.nn as nn
sys.path.insert(0, sys.argv[1])
from multimodal_gar_b200 import train_step as TS

class StubBackbone(nn.Module):
    # same contract as PointLiDARBackbone: data_dict -> (B, MNP, dim); per-person statistics of the in-box-ish points
    def __init__(self, dim):
        super().__init__()
        self.embedding = nn.Linear(8, dim)
        self.bn = nn.BatchNorm1d(8)                       # batch statistics: per rank, like DataParallel replicas
    def forward(self, d):
        pts, boxes = d['points'], d['gt_bboxes3d']
        ctr = boxes[..., :3]
        near = (pts[:, None, :, :3] - ctr[:, :, None, :]).abs().amax(-1) < 1.0          # (B, MNP, N)
        w = near.float()
        cnt = w.sum(-1, keepdim=True).clamp_min(1.0)
        mean = (w.unsqueeze(-1) * pts[:, None]).sum(2) / cnt                              # (B, MNP, 4)
        f = torch.cat([mean, mean * mean], dim=-1)
        B, P, C = f.shape
        return self.embedding(self.bn(f.view(B * P, C)).view(B, P, C))

dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
torch.manual_seed(3)
dim = 32
net = TS.LidarFusionNet(dim=dim, backbone=StubBackbone(dim))
twin = copy.deepcopy(net)
full = TS.synthetic_batch(11, 4, N=512, M=6, MNP=8, dim=dim)
mine = TS.shard_batch(full, world, rank)
ddp = TS.make_ddp(net)
opt = torch.optim.Adam(ddp.parameters(), lr=1e-3)
loss = ddp(mine)
loss.backward()
grads = [p.grad.clone() for p in net.parameters()]
# (1) every rank holds the same gradients after the all-reduce
for g in grads:
    other = [torch.zeros_like(g) for _ in range(world)]
    dist.all_gather(other, g)
    assert all(torch.equal(o, other[0]) for o in other)
# (2) they are the mean of the per-shard gradients of one process
want = [torch.zeros_like(p) for p in twin.parameters()]
for r in range(world):
    t = copy.deepcopy(twin)
    t(TS.shard_batch(full, world, r)).backward()
    for w_, p in zip(want, t.parameters()):
        w_ += p.grad / world
for (name, _), g, w_ in zip(net.named_parameters(), grads, want):
    scale = max(1e-6, float(w_.abs().max()))
    assert torch.allclose(g, w_, rtol=1e-5, atol=1e-5 * scale), (name, float((g - w_).abs().max()) / scale)
# (3) the replicas stay identical over optimizer steps; BatchNorm running statistics stay per rank (not broadcast)
for step in range(2):
    opt.step()
    TS.train_step(ddp, opt, mine)
for p in net.parameters():
    other = [torch.zeros_like(p) for _ in range(world)]
    dist.all_gather(other, p.detach())
    assert all(torch.equal(o, other[0]) for o in other)
rm = net.lidar_backbone.bn.running_mean
other = [torch.zeros_like(rm) for _ in range(world)]
dist.all_gather(other, rm)
assert not torch.equal(other[0], other[1]), "running statistics should follow each rank's own frames"
assert TS.grad_bytes(net) == 4 * sum(p.numel() for p in net.parameters())
if rank == 0:
    print("DDP_OK", world, float(loss))
dist.barrier()
dist.destroy_process_group()
"""


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_ddp_gradients_match_single_process_over_gloo_world2(tmp_path):
    script = tmp_path / "ddp_worker.py"
    script.write_text(_WORKER)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), str(script), ROOT]
    out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-4000:]
    assert "DDP_OK 2" in out.stdout


def test_shard_batch_partitions_every_key():
    from multimodal_gar_b200 import train_step as TS
    full = TS.synthetic_batch(5, 5, N=64, M=3, MNP=4, dim=8)
    parts = [TS.shard_batch(full, 2, r) for r in range(2)]
    for k, v in full.items():
        assert torch.equal(torch.cat([p[k] for p in parts]), v), k
    assert parts[0]['points'].shape[0] + parts[1]['points'].shape[0] == 5


@pytest.mark.gpu
def test_training_step_through_the_cuda_branch(cuda):
    """One rank of the configs[3] step with the real LiDAR branch: finite loss that decreases over a few Adam steps,
    gradients on every parameter (SA-MLP weights included: the train kernels ran), and the single-rank
    DistributedDataParallel wrapper (NCCL, world 1) produces the same gradients as the bare module."""
    import copy

    import torch.distributed as dist

    from multimodal_gar_b200 import _lib, train_step as TS
    torch.manual_seed(5)
    net = TS.LidarFusionNet().to(cuda).train()
    batch = TS.synthetic_batch(21, 2, N=16384, M=12, MNP=16, device=cuda)
    bare = copy.deepcopy(net)
    n0 = _lib.launch_count()
    bare(batch).backward()
    assert _lib.launch_count() - n0 >= 10, "the CUDA library did not run the training kernels"
    for name, p in bare.named_parameters():
        assert p.grad is not None and torch.isfinite(p.grad).all(), name
    assert float(bare.lidar_backbone.encoder.sa.mlps[0][0].weight.grad.abs().sum()) > 0
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", str(_free_port()))
    created = not dist.is_initialized()
    if created:
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=cuda)
    try:
        ddp = TS.make_ddp(net, cuda)
        ddp(batch).backward()
        for (name, a), b in zip(net.named_parameters(), bare.parameters()):
            scale = max(1e-6, float(b.grad.abs().max()))
            assert torch.allclose(a.grad, b.grad, rtol=1e-4, atol=1e-4 * scale), name     # float atomics reorder sums
        opt = torch.optim.Adam(ddp.parameters(), lr=1e-3)
        losses = [float(TS.train_step(ddp, opt, batch)) for _ in range(6)]
        assert all(l == l for l in losses) and losses[-1] < losses[0]
    finally:
        if created:
            dist.destroy_process_group()
