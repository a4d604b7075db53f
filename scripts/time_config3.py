"""configs[2] at full size on one GPU: 64 panoramic frames x 80 persons (zero-padded to 100 proposals, 65 536 points per
frame): full LiDAR branch (PointLiDARBackbone -> (B,100,512)) + DAFM input path + DistanceAwareFusion forward on synthetic
RGB features, CUDA events; the fusion step with the fused kernels (csrc/dafm.cu) and with the torch composition."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from multimodal_gar_b200 import dafm, lidar_backbone as LB, synthetic  # noqa: E402
from scripts.time_dafm import timed  # noqa: E402


def main():
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
    bboxes = torch.cat([bb2, bb2 + 20 + 60 * torch.rand((B, MNP, 2), device=dev)], dim=-1)
    pid = torch.full((B, MNP), -1, dtype=torch.int64, device=dev)
    pid[:, :M] = torch.arange(M, device=dev)
    res = {"workload": "configs[2]: %d frames x %d persons (padded to %d), %d points per frame, LiDAR branch + DAFM forward" % (B, M, MNP, N)}
    with torch.no_grad():
        lidar = net(dict(points=points, gt_bboxes3d=padded))
        d = LB.dafm_lidar_inputs(lidar, padded, pid)
        res["lidar_branch_ms"] = timed(lambda: net(dict(points=points, gt_bboxes3d=padded)), iters=10)
        res["dafm_inputs_ms"] = timed(lambda: LB.dafm_lidar_inputs(lidar, padded, pid), iters=10)
        for name in ("fused", "torch"):
            os.environ["MGAR_DAFM"] = name
            res["dafm_%s_ms" % name] = timed(lambda: fus(rgb, d["feature"], bboxes, padded, d["valid"]), iters=10)

            def whole():
                f = net(dict(points=points, gt_bboxes3d=padded))
                dd = LB.dafm_lidar_inputs(f, padded, pid)
                return fus(rgb, dd["feature"], bboxes, padded, dd["valid"])
            res["total_%s_ms" % name] = timed(whole, iters=10)
            res["frames_per_s_%s" % name] = round(B / res["total_%s_ms" % name] * 1e3, 1)
    print(json.dumps(res))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/config3_timing.json", "w"))


if __name__ == "__main__":
    main()
