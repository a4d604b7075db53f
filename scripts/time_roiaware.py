"""RoIAwarePool3d forward/backward: the new kernels and the unmodified reference extension timed on the same
inputs with CUDA events (one 65536-point frame, 40 person boxes, 128 channels, 12^3 voxels, 128 points per voxel -
the reference's defaults in pcdet's PartA2 head)."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from multimodal_gar_b200 import synthetic  # noqa: E402
from multimodal_gar_b200.pcdet.ops.roiaware_pool3d import roiaware_pool3d_cuda as new  # noqa: E402
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
    return a.elapsed_time(b) / iters


def main():
    dev = torch.device("cuda:0")
    ref = O.load_ref("roiaware_pool3d_cuda")
    N, M, C, out, mp = 65536, 40, 128, 12, 128
    pts, boxes = synthetic.make_frame(1994, N=N, M=M)
    p = torch.from_numpy(np.ascontiguousarray(pts[:, :3])).to(dev)
    b = torch.from_numpy(boxes).to(dev)
    f = torch.randn(N, C, device=dev)
    res = {"workload": "N=%d M=%d C=%d out=%d^3 max_pts=%d" % (N, M, C, out, mp)}
    for name, mod in (("new", new), ("reference", ref)):
        if mod is None:
            continue
        for mi, method in enumerate(("max", "avg")):
            pooled = torch.zeros(M, out, out, out, C, device=dev)
            argmax = torch.zeros(M, out, out, out, C, dtype=torch.int32, device=dev)
            lists = torch.zeros(M, out, out, out, mp, dtype=torch.int32, device=dev)

            def fwd():
                lists.zero_()
                mod.forward(b, p, f, argmax, lists, pooled, mi)
            res["%s_fwd_%s_ms" % (name, method)] = round(timed(fwd), 4)
            g = torch.randn_like(pooled)
            gin = torch.zeros(N, C, device=dev)
            res["%s_bwd_%s_ms" % (name, method)] = round(timed(lambda: mod.backward(lists, argmax, g, gin, mi)), 4)
    print(json.dumps(res))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/roiaware_timing.json", "w"))


if __name__ == "__main__":
    main()
