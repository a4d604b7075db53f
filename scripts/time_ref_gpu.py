"""the reference's own CUDA extensions (oracle/_ref, unmodified sources) through the same configs[1]
pipeline on the same B200, timed beside the new path - the "reference-GPU" row.  Measurement infrastructure like
the tests (it executes oracle/_ref); not part of bench.py.    python scripts/time_ref_gpu.py [frames] > out.json
Composition = what the reference's Python does: points_in_boxes_gpu, RoIPointPool3d (kernels + its two cudaMalloc),
canonical transform (torch), farthest_point_sample, QueryAndGroup (ball query + 2 groupings + torch glue),
[Conv2d, BatchNorm2d, ReLU] x 2 in eval mode (cuDNN, TF32 allowed as by torch's default for convolutions), max pools."""
import json
import os
import sys

import numpy as np
import torch
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from multimodal_gar_b200 import lidar_branch  # noqa: E402
from oracle import oracle as O  # noqa: E402


def measure(F=16, dev=None):
    """-> dict: per-stage and total ms of the reference composition on F configs[1] frames (current CUDA device)."""
    dev = dev or torch.device("cuda", torch.cuda.current_device())
    roiaware, roipoint, pn2 = (O.load_ref(n) for n in ("roiaware_pool3d_cuda", "roipoint_pool3d_cuda", "pointnet2_stack_cuda"))
    assert roiaware and roipoint and pn2, "oracle/_ref is not built"
    c = bench.CFG
    N, M, S, P, K, r = c["N"], c["M"], c["S"], c["P"], c["K"], c["radius"]
    xyz, feat, boxes = (torch.from_numpy(a).to(dev) for a in bench.make_frames(F))
    enc = bench.make_encoder().to(dev)
    mlp = enc.sa.mlps[0]

    def ref_step(marks=None):
        def mark(n):
            if marks is not None:
                e = torch.cuda.Event(enable_timing=True); e.record(); marks.append((n, e))
        mark("start")
        box_idx = torch.full((F, N), -1, dtype=torch.int32, device=dev)
        roiaware.points_in_boxes_gpu(boxes, xyz, box_idx)
        mark("points_in_boxes")
        pooled = torch.zeros((F, M, S, 4), device=dev)
        empty = torch.zeros((F, M), dtype=torch.int32, device=dev)
        roipoint.forward(xyz, boxes, feat, pooled, empty)
        pooled = lidar_branch.canonical_transform(pooled, boxes, empty)
        mark("roipoint_pool")
        cx = pooled[:, :, 0:3].contiguous()
        temp = torch.full((F * M, S), 1e10, device=dev)
        idx = torch.zeros((F * M, P), dtype=torch.int32, device=dev)
        pn2.farthest_point_sampling_wrapper(F * M, S, P, cx, temp, idx)
        mark("fps")
        new_xyz = torch.gather(cx, 1, idx.long().unsqueeze(-1).expand(-1, -1, 3)).view(-1, 3)
        flat = cx.view(-1, 3)
        cnt = torch.full((F * M,), S, dtype=torch.int32, device=dev)
        ncnt = torch.full((F * M,), P, dtype=torch.int32, device=dev)
        bidx = torch.zeros((F * M * P, K), dtype=torch.int32, device=dev)
        pn2.ball_query_wrapper(F * M, F * M * P, r, K, new_xyz, ncnt, flat, cnt, bidx)
        emptyb = bidx[:, 0] == -1
        bidx[emptyb] = 0
        mark("ball_query")
        g_xyz = torch.zeros((F * M * P, 3, K), device=dev)
        pn2.group_points_wrapper(F * M, F * M * P, 3, K, flat, cnt, bidx, ncnt, g_xyz)
        g_xyz -= new_xyz.unsqueeze(-1)
        g_xyz[emptyb] = 0
        pf = pooled[:, :, 3:].contiguous().view(-1, 1)
        g_f = torch.zeros((F * M * P, 1, K), device=dev)
        pn2.group_points_wrapper(F * M, F * M * P, 1, K, pf, cnt, bidx, ncnt, g_f)
        g_f[emptyb] = 0
        nf = torch.cat([g_xyz, g_f], dim=1).permute(1, 0, 2).unsqueeze(0)
        out = mlp(nf)
        out = nn.functional.max_pool2d(out, kernel_size=[1, K]).squeeze(-1).squeeze(0).permute(1, 0)
        person = out.reshape(F * M, P, -1).max(dim=1).values * (empty.view(-1, 1) == 0)
        mark("group+sa_mlp")
        return person

    with torch.no_grad():
        for _ in range(2):
            ref_step()
        torch.cuda.synchronize()
        marks = []
        steps = 3
        for _ in range(steps):
            ref_step(marks)
        torch.cuda.synchronize()
        stage = {}
        for (n0, e0), (n1, e1) in zip(marks, marks[1:]):
            if n1 != "start":
                stage[n1] = stage.get(n1, 0.0) + e0.elapsed_time(e1) / steps
        ms = sum(stage.values())
        new = enc(xyz, feat, boxes)
        ref = ref_step().view(F, M, -1)
        scale = float(ref.abs().max())
        diff = float((new - ref).abs().max())
    return dict(what="reference CUDA extensions (oracle/_ref) + torch/cuDNN composition, configs[1], eval mode",
                frames_per_step=F, ms_per_step=ms, frames_per_s=F / ms * 1e3,
                stage_ms={k: round(v, 3) for k, v in stage.items()},
                max_abs_diff_vs_new_path=diff, feature_scale=scale,
                note="cuDNN runs the two 1x1 convolutions in TF32 (torch default), so the difference to the "
                     "fp32-faithful new path is TF32-sized")


def main():
    print(json.dumps(measure(int(sys.argv[1]) if len(sys.argv) > 1 else 16)))


if __name__ == "__main__":
    main()
