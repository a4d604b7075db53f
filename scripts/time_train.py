"""Training step of the SA module at hot-path size (F frames x 40 persons x 512 points, 512 centroids, 32 samples,
MLP 4-64-128): forward with batch-statistics BatchNorm + full backward, native kernels (sa_mlp_train.cu) against the
reference's composition (grouping kernel + Conv2d/BatchNorm2d/ReLU/max_pool under torch autograd) on the same GPU."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_stack import pointnet2_modules  # noqa: E402


def timed(fn, iters=5):
    for _ in range(2):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return round(a.elapsed_time(b) / iters, 3)


def main():
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    F_, M, S, P = int(os.environ.get("FRAMES", 8)), 40, 512, 512
    B = F_ * M
    xyz = (torch.rand(B * S, 3, device=dev) - 0.5) * torch.tensor([0.8, 0.8, 1.8], device=dev)
    feats = torch.rand(B * S, 1, device=dev)
    cnt = torch.full((B,), S, dtype=torch.int32, device=dev)
    new_xyz = xyz.view(B, S, 3)[:, :P].reshape(-1, 3).contiguous()
    ncnt = torch.full((B,), P, dtype=torch.int32, device=dev)
    mod = pointnet2_modules.StackSAModuleMSG(radii=[0.3], nsamples=[32], mlps=[[1, 64, 128]], use_xyz=True,
                                             pool_method="max_pool").to(dev).train()
    res = {"workload": "%d frames x 40 persons: %d centroids x 32 samples, MLP 4-64-128, train mode fwd+bwd" % (F_, B * P)}

    def step():
        f = feats.clone().requires_grad_(True)
        _, out = mod(xyz, cnt, new_xyz, ncnt, f)
        out.square().sum().backward()
        mod.zero_grad(set_to_none=True)

    for name in os.environ.get("MODES", "native,native_tc,native_generic,torch").split(","):
        os.environ["MGAR_SA_TRAIN"] = "torch" if name == "torch" else "native"
        os.environ["MGAR_SA_TRAIN_KERNELS"] = {"native_generic": "generic", "native_tc": "tc"}.get(name, "fast")
        if name == "native_generic" and F_ > 16:
            continue
        torch.cuda.reset_peak_memory_stats()
        res[name + "_ms"] = timed(step)
        res[name + "_peak_GB"] = round(torch.cuda.max_memory_allocated() / 2 ** 30, 2)
        res["frames_per_s_" + name] = round(F_ / res[name + "_ms"] * 1e3, 1)
    print(json.dumps(res))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/train_timing.json", "w"))


if __name__ == "__main__":
    main()
