"""configs[2]'s fusion step: DistanceAwareFusion over 64 frames x 100 padded persons (80 real), D = 512 - the fused path
(csrc/dafm.cu between the library GEMMs) against the torch composition of the same module, CUDA events."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from multimodal_gar_b200 import dafm  # noqa: E402


def timed(fn, iters=30):
    for _ in range(5):
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
    B, n, real, D = 64, 100, 80, 512
    fus = dafm.DistanceAwareFusion(D).to(dev).eval()
    rgb, lidar = torch.randn(B, n, D, device=dev), torch.randn(B, n, D, device=dev)
    b2 = torch.rand(B, n, 4, device=dev)
    b2[..., 2:] += b2[..., :2] + 0.1
    b3 = torch.rand(B, n, 7, device=dev) * 15
    valid = (torch.arange(n, device=dev).view(1, -1) < real).expand(B, n).contiguous()
    res = {"workload": "DistanceAwareFusion (2 x FusionAttention_mat), %d frames x %d persons (%d real), D=%d" % (B, n, real, D)}
    with torch.no_grad():
        for name in ("fused", "torch"):
            os.environ["MGAR_DAFM"] = name
            res[name + "_ms"] = timed(lambda: fus(rgb, lidar, b2, b3, valid))
            res["frames_per_s_" + name] = round(B / res[name + "_ms"] * 1e3, 1)
    res["gemm_flops"] = 2 * 2 * 10 * B * n * D * D
    print(json.dumps(res))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/dafm_timing.json", "w"))


if __name__ == "__main__":
    main()
