"""Voxel query: the new kernel and the unmodified reference extension timed on the same inputs with CUDA events
(VoxelRCNN-like shape: 2 frames, 41x100x88 grid, 60k occupied voxels per frame, 128 RoIs x 216 grid points per frame,
range [4,4,4], 16 samples)."""
import json
import os
import sys

import torch

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases  # noqa: E402
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
    return a.elapsed_time(b) / iters


def main():
    dev = torch.device("cuda:0")
    sc = cases.voxel_scene(seed=8, B=2, grid=(41, 100, 88), voxels=60000, queries_per_frame=27648, C=1)
    t = {k: torch.from_numpy(sc[k]).to(dev) for k in ("xyz", "new_xyz", "new_coords", "v2p")}
    M = sc["new_coords"].shape[0]
    res = {"workload": "M=%d queries, grid 41x100x88 x2, range 4,4,4, nsample 16" % M}
    for radius in (0.8, 0.2):
        for name, mod in (("new", new), ("reference", O.load_ref("pointnet2_stack_cuda"))):
            if mod is None:
                continue
            idx = torch.zeros((M, 16), dtype=torch.int32, device=dev)
            res["%s_r%.1f_ms" % (name, radius)] = round(timed(lambda: mod.voxel_query_wrapper(
                M, 41, 100, 88, 16, radius, 4, 4, 4, t["new_xyz"], t["xyz"], t["new_coords"], t["v2p"], idx)), 4)
    print(json.dumps(res))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/voxel_query_timing.json", "w"))


if __name__ == "__main__":
    main()
