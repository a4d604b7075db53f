"""Generates tests/golden/ref_voxel_query.npz by running the UNMODIFIED reference extension
(oracle/_ref/pointnet2_stack_cuda.so, compiled from /root/reference/pcdet/ops/pointnet2/pointnet2_stack/src) on a GPU:

    gpurun -- python tests/golden/make_ref_voxel_query_golden.py     # writes gpurun_out/ref_voxel_query.npz
    cp gpurun_out/ref_voxel_query.npz tests/golden/

The CPU suite (tests/test_oracle_pin.py) pins the C oracle of the voxel query to these rows; the scenes are rebuilt
from their seeds (tests/cases.py::voxel_scene), only the kernel outputs are stored."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(HERE, ".."))
import cases  # noqa: E402
from oracle import oracle as O  # noqa: E402

SCENES = [dict(seed=5, B=2, grid=(6, 24, 24), voxels=1500, queries_per_frame=90, rng=(2, 2, 2), radius=0.3, nsample=8),
          dict(seed=6, B=3, grid=(10, 60, 60), voxels=20000, queries_per_frame=400, rng=(4, 4, 4), radius=0.45, nsample=16),
          dict(seed=7, B=1, grid=(3, 9, 200), voxels=3000, queries_per_frame=301, rng=(0, 1, 5), radius=10.0, nsample=40)]


def main():
    ref = O.load_ref("pointnet2_stack_cuda")
    dev = torch.device("cuda:0")
    out = {}
    for i, case in enumerate(SCENES):
        c = dict(case)
        rng, radius, nsample = c.pop("rng"), c.pop("radius"), c.pop("nsample")
        sc = cases.voxel_scene(C=1, **c)
        t = {k: torch.from_numpy(sc[k]).to(dev) for k in ("xyz", "new_xyz", "new_coords", "v2p")}
        M = sc["new_coords"].shape[0]
        _, Z, Y, X = sc["v2p"].shape
        idx = torch.zeros((M, nsample), dtype=torch.int32, device=dev)
        ref.voxel_query_wrapper(M, Z, Y, X, nsample, radius, rng[0], rng[1], rng[2], t["new_xyz"], t["xyz"], t["new_coords"],
                                t["v2p"], idx)
        out["idx_%d" % i] = idx.cpu().numpy()
        same = np.array_equal(out["idx_%d" % i], O.voxel_query_raw(rng, radius, nsample, sc["xyz"], sc["new_xyz"],
                                                                   sc["new_coords"], sc["v2p"]))
        print("scene", i, "rows", M, "empty", int((out["idx_%d" % i][:, 0] == -1).sum()), "oracle == reference:", same)
    os.makedirs("gpurun_out", exist_ok=True)
    np.savez_compressed("gpurun_out/ref_voxel_query.npz", **out)
    print("wrote gpurun_out/ref_voxel_query.npz")


if __name__ == "__main__":
    main()
