"""pointnet2_batch surface: the new kernels and the unmodified reference extension timed on the same inputs with
CUDA events (PointNet2MSG-like: 4 frames x 16384 points, 4096 centroids, 32 samples, 64 channels)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_batch import pointnet2_batch_cuda as new  # noqa: E402
from oracle import oracle as O  # noqa: E402


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return round(a.elapsed_time(b) / iters, 4)


def main():
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    B, N, M, ns, C = 4, 16384, 4096, 32, 64
    xyz = torch.rand(B, N, 3, device=dev) * torch.tensor([40.0, 40.0, 3.0], device=dev)
    new_xyz = xyz[:, :M].contiguous()
    feats = torch.randn(B, C, N, device=dev)
    res = {"workload": "B=%d N=%d M=%d nsample=%d C=%d" % (B, N, M, ns, C)}
    for name, mod in (("new", new), ("reference", O.load_ref("pointnet2_batch_cuda"))):
        if mod is None:
            continue
        idx = torch.zeros(B, M, ns, dtype=torch.int32, device=dev)
        res[name + "_ball_query_ms"] = timed(lambda: mod.ball_query_wrapper(B, N, M, 1.0, ns, new_xyz, xyz, idx))
        out = torch.empty(B, C, M, ns, device=dev)
        res[name + "_group_ms"] = timed(lambda: mod.group_points_wrapper(B, C, N, M, ns, feats, idx, out))
        g = torch.zeros_like(feats)
        res[name + "_group_grad_ms"] = timed(lambda: mod.group_points_grad_wrapper(B, C, N, M, ns, out, idx, g))
        d2 = torch.empty(B, N, 3, device=dev)
        i3 = torch.empty(B, N, 3, dtype=torch.int32, device=dev)
        res[name + "_three_nn_ms"] = timed(lambda: mod.three_nn_wrapper(B, N, M, xyz, new_xyz, d2, i3))
        kf = torch.randn(B, C, M, device=dev)
        w = torch.rand(B, N, 3, device=dev)
        up = torch.empty(B, C, N, device=dev)
        res[name + "_interpolate_ms"] = timed(lambda: mod.three_interpolate_wrapper(B, C, M, N, kf, i3, w, up))
        gk = torch.zeros_like(kf)
        res[name + "_interpolate_grad_ms"] = timed(lambda: mod.three_interpolate_grad_wrapper(B, C, N, M, up, i3, w, gk))
    print(json.dumps(res))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/batch_ops_timing.json", "w"))


if __name__ == "__main__":
    main()
