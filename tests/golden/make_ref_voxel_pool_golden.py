"""Golden vectors for NeighborVoxelSAModuleMSG, produced by the REFERENCE class itself.

The class is cut out of /root/reference/pcdet/ops/pointnet2/pointnet2_stack/voxel_pool_modules.py by its source
lines and executed unmodified on CPU tensors; the one thing it cannot do here is call the CUDA voxel query /
grouping (no GPU in the build container), so `voxel_query_utils.VoxelQueryAndGrouping` is a stand-in that returns the
CPU oracle's result (oracle.voxel_query_and_group, itself pinned to the reference kernels by
tests/golden/ref_voxel_query.npz).  eval() and train() mode (batch statistics) outputs are stored.
Run in the build container only:   python tests/golden/make_ref_voxel_pool_golden.py  -> tests/golden/ref_voxel_pool.npz
"""
import ast
import os
import sys
import types

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(HERE, ".."))
import cases  # noqa: E402
from oracle import oracle as O  # noqa: E402

SRC = "/root/reference/pcdet/ops/pointnet2/pointnet2_stack/voxel_pool_modules.py"


class CpuVoxelQueryAndGrouping(nn.Module):
    def __init__(self, max_range, radius, nsample):
        super().__init__()
        self.max_range, self.radius, self.nsample = max_range, radius, nsample

    def forward(self, new_coords, xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt, features, voxel2point_indices):
        gf, gx, empty = O.voxel_query_and_group(self.max_range, self.radius, self.nsample, new_coords.numpy(), xyz.numpy(),
                                                xyz_batch_cnt.numpy(), new_xyz.numpy(), new_xyz_batch_cnt.numpy(),
                                                features.detach().numpy(), voxel2point_indices.numpy())
        return torch.from_numpy(gf), torch.from_numpy(gx), torch.from_numpy(empty)


def reference_class():
    text = open(SRC).read()
    node = [n for n in ast.parse(text).body if isinstance(n, ast.ClassDef) and n.name == "NeighborVoxelSAModuleMSG"][0]
    code = "\n".join(text.splitlines()[node.lineno - 1:node.end_lineno])
    from typing import List
    vq = types.SimpleNamespace(VoxelQueryAndGrouping=CpuVoxelQueryAndGrouping)
    ns = dict(torch=torch, nn=nn, F=F, List=List, voxel_query_utils=vq)
    exec(compile(code, SRC, "exec"), ns)
    return ns["NeighborVoxelSAModuleMSG"]


def main():
    torch.manual_seed(1994)
    Ref = reference_class()
    sc = cases.voxel_scene(seed=5, B=2, grid=(6, 24, 24), voxels=1500, queries_per_frame=90, C=16)
    kw = dict(query_ranges=[[2, 2, 2], [1, 3, 3]], radii=[0.3, 0.45], nsamples=[8, 16], mlps=[[16, 16, 24], [16, 32, 8]])
    out = dict(sc)
    for pool in ("max_pool", "avg_pool"):
        torch.manual_seed(7)
        m = Ref(pool_method=pool, **kw)
        with torch.no_grad():
            for mod in m.modules():
                if isinstance(mod, (nn.BatchNorm1d, nn.BatchNorm2d)):
                    mod.running_mean.normal_(0, 0.1)
                    mod.running_var.uniform_(0.8, 1.2)
                    mod.weight.normal_(1, 0.1)
                    mod.bias.normal_(0, 0.1)
        for k, v in m.state_dict().items():
            out["%s.%s" % (pool, k)] = v.numpy().copy()
        t = {k: torch.from_numpy(v) for k, v in sc.items()}
        args = (t["xyz"], t["xyz_batch_cnt"], t["new_xyz"], t["new_xyz_batch_cnt"], t["new_coords_xyz"], t["features"], t["v2p"])
        with torch.no_grad():
            m.eval()
            out[pool + ".eval"] = m(*args).numpy()
            m.train()
            out[pool + ".train"] = m(*args).numpy()
    np.savez_compressed(os.path.join(HERE, "ref_voxel_pool.npz"), **out)
    print("wrote ref_voxel_pool.npz", out["max_pool.eval"].shape, float(np.abs(out["max_pool.eval"]).mean()))


if __name__ == "__main__":
    main()
