"""ONE configs[2] step (64 frames x 80 persons padded to 100: LiDAR branch + DAFM inputs + fused DAFM) after two warm-up
steps, for an ncu launch list:  ncu --metrics gpu__time_duration.sum --clock-control none --csv ... python scripts/one_config3_step.py
(cudaProfilerStart/Stop bracket the step: run ncu with --profile-from-start off)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from multimodal_gar_b200 import dafm, lidar_backbone as LB, synthetic  # noqa: E402

dev = torch.device("cuda:0")
B, M, MNP, N = 64, 80, 100, 65536
pts, boxes = synthetic.make_batch(1994, B, N=N, M=M)
padded = torch.zeros((B, MNP, 7), device=dev)
padded[:, :M] = torch.from_numpy(boxes).to(dev)
points = torch.from_numpy(np.ascontiguousarray(pts)).to(dev)
torch.manual_seed(9)
net = LB.PointLiDARBackbone().to(dev).eval()
fus = dafm.DistanceAwareFusion(512).to(dev).eval()
rgb = torch.randn((B, MNP, 512), device=dev)
bb2 = torch.rand((B, MNP, 2), device=dev) * 500
bboxes = torch.cat([bb2, bb2 + 80], dim=-1)
pid = torch.full((B, MNP), -1, dtype=torch.int64, device=dev)
pid[:, :M] = torch.arange(M, device=dev)


def step():
    f = net(dict(points=points, gt_bboxes3d=padded))
    d = LB.dafm_lidar_inputs(f, padded, pid)
    return fus(rgb, d["feature"], bboxes, padded, d["valid"])


with torch.no_grad():
    step(); step()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    out = step()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
print("ok", tuple(out.shape))
