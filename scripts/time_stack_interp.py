"""Stack-layout feature propagation: the new kernels and the unmodified reference extension timed on the same inputs
with CUDA events (4 clouds x 16384 unknown points, 4096 known points each, 128 channels)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_stack import pointnet2_stack_cuda as new  # noqa: E402
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
    B, N, M, C = 4, 16384, 4096, 128
    unknown = torch.rand(B * N, 3, device=dev) * torch.tensor([40.0, 40.0, 3.0], device=dev)
    known = unknown.view(B, N, 3)[:, :M].reshape(-1, 3).contiguous()
    ucnt = torch.full((B,), N, dtype=torch.int32, device=dev)
    kcnt = torch.full((B,), M, dtype=torch.int32, device=dev)
    feats = torch.randn(B * M, C, device=dev)
    w = torch.rand(B * N, 3, device=dev)
    res = {"workload": "B=%d N=%d M=%d C=%d (stack layout)" % (B, N, M, C)}
    for name, mod in (("new", new), ("reference", O.load_ref("pointnet2_stack_cuda"))):
        if mod is None:
            continue
        d2 = torch.empty(B * N, 3, device=dev)
        i3 = torch.empty(B * N, 3, dtype=torch.int32, device=dev)
        res[name + "_three_nn_ms"] = timed(lambda: mod.three_nn_wrapper(unknown, ucnt, known, kcnt, d2, i3))
        up = torch.empty(B * N, C, device=dev)
        res[name + "_interpolate_ms"] = timed(lambda: mod.three_interpolate_wrapper(feats, i3, w, up))
        gk = torch.zeros_like(feats)
        res[name + "_interpolate_grad_ms"] = timed(lambda: mod.three_interpolate_grad_wrapper(up, i3, w, gk))
    res["interpolate_algorithmic_GBps_new"] = round((B * N * C * 4 * 2 + B * N * 24) / res["new_interpolate_ms"] / 1e6, 1)
    print(json.dumps(res))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/stack_interp_timing.json", "w"))


if __name__ == "__main__":
    main()
