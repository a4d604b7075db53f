"""Generates tests/golden/ref_roiaware.npz by running the UNMODIFIED reference extension
(oracle/_ref/roiaware_pool3d_cuda.so, compiled from /root/reference/pcdet/ops/roiaware_pool3d/src) on a GPU:

    gpurun -- python tests/golden/make_ref_roiaware_golden.py      # writes gpurun_out/ref_roiaware.npz
    cp gpurun_out/ref_roiaware.npz tests/golden/

The CPU suite (tests/test_oracle_pin.py) then pins the C oracle of RoIAwarePool3d to these outputs.  Seeds are
searched for a case where host and device trigonometry agree on every point (no voxel-face ambiguity), so the
CPU comparison can be exact."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from multimodal_gar_b200 import synthetic  # noqa: E402
from oracle import oracle as O  # noqa: E402


def main():
    ref = O.load_ref("roiaware_pool3d_cuda")
    dev = torch.device("cuda:0")
    out = {}
    N, M, C, size, max_pts = 4096, 6, 4, (3, 4, 2), 6
    for seed in range(100, 140):
        pts, boxes = synthetic.make_frame(seed, N=N, M=M, density_scale=4.0)
        pts = np.ascontiguousarray(pts[:, :3])
        feat = np.random.default_rng(seed).standard_normal((N, C)).astype(np.float32)
        ok, rec = True, {}
        for mi, method in enumerate(("max", "avg")):
            t = [torch.from_numpy(a).to(dev) for a in (boxes, pts, feat)]
            pooled = torch.zeros((M,) + size + (C,), device=dev)
            argmax = torch.zeros((M,) + size + (C,), dtype=torch.int32, device=dev)
            lists = torch.zeros((M,) + size + (max_pts,), dtype=torch.int32, device=dev)
            ref.forward(t[0], t[1], t[2], argmax, lists, pooled, mi)
            g = torch.from_numpy(np.random.default_rng(seed + mi).standard_normal(pooled.shape).astype(np.float32)).to(dev)
            gin = torch.zeros((N, C), device=dev)
            ref.backward(lists, argmax, g, gin, mi)
            torch.cuda.synchronize()
            w = O.roiaware_pool3d(boxes, pts, feat, size, max_pts, method)
            ok &= all(np.array_equal(a, b.cpu().numpy()) for a, b in zip(w, (pooled, argmax, lists)))
            rec.update({method + "_pooled": pooled.cpu().numpy(), method + "_argmax": argmax.cpu().numpy(),
                        method + "_lists": lists.cpu().numpy(), method + "_grad_out": g.cpu().numpy(),
                        method + "_grad_in": gin.cpu().numpy()})
        print("seed", seed, "oracle == reference:", ok, "hits", int(rec["max_lists"][..., 0].sum()))
        if ok and rec["max_lists"][..., 0].sum() > 100:
            out = dict(rec, boxes=boxes, pts=pts, feat=feat, size=np.array(size), max_pts=np.array(max_pts), seed=np.array(seed))
            break
    assert out, "no seed without host/device trig ambiguity found"
    os.makedirs("gpurun_out", exist_ok=True)
    np.savez_compressed("gpurun_out/ref_roiaware.npz", **out)
    print("wrote gpurun_out/ref_roiaware.npz")


if __name__ == "__main__":
    main()
