"""Golden vectors for PointnetSAModuleMSG / PointnetFPModule, produced by the REFERENCE classes.

The classes are cut out of /root/reference/pcdet/ops/pointnet2/pointnet2_batch/pointnet2_modules.py by their source
lines and executed unmodified on CPU tensors; the CUDA functions they call through `pointnet2_utils` (FPS, gather,
ball query + grouping, three_nn, three_interpolate - no GPU in the build container) are stand-ins that return the CPU
oracle's results, which tests/test_batch_ops_gpu.py pins to the reference kernels on the B200.
Run in the build container only:  python tests/golden/make_ref_batch_modules_golden.py -> tests/golden/ref_batch_modules.npz
"""
import ast
import os
import sys
import types
from typing import List

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import oracle as O  # noqa: E402

SRC = "/root/reference/pcdet/ops/pointnet2/pointnet2_batch/pointnet2_modules.py"
T = torch.from_numpy


class CpuQueryAndGroup(nn.Module):
    def __init__(self, radius, nsample, use_xyz=True):
        super().__init__()
        self.radius, self.nsample, self.use_xyz = radius, nsample, use_xyz

    def forward(self, xyz, new_xyz, features=None):                      # pointnet2_utils.py:236-258 on the oracle
        idx = O.batch_ball_query(self.radius, self.nsample, xyz.numpy(), new_xyz.numpy())
        g = T(O.batch_grouping_operation(xyz.transpose(1, 2).contiguous().numpy(), idx))
        g = g - new_xyz.transpose(1, 2).unsqueeze(-1)
        if features is None:
            return g
        gf = T(O.batch_grouping_operation(features.contiguous().numpy(), idx))
        return torch.cat([g, gf], dim=1) if self.use_xyz else gf


class CpuGroupAll(nn.Module):
    def __init__(self, use_xyz=True):
        super().__init__()
        self.use_xyz = use_xyz

    def forward(self, xyz, new_xyz, features=None):
        g = xyz.transpose(1, 2).unsqueeze(2)
        if features is None:
            return g
        return torch.cat([g, features.unsqueeze(2)], dim=1) if self.use_xyz else features.unsqueeze(2)


def _three_nn(unknown, known):
    d, i, _ = O.three_nn(unknown.numpy(), known.numpy())
    return T(d), T(i)


utils = types.SimpleNamespace(
    QueryAndGroup=CpuQueryAndGroup, GroupAll=CpuGroupAll,
    farthest_point_sample=lambda xyz, n: T(O.farthest_point_sample(xyz.numpy(), n)),
    gather_operation=lambda f, idx: T(O.batch_grouping_operation(f.numpy(), idx.numpy())),
    three_nn=_three_nn,
    three_interpolate=lambda f, i, w: T(O.three_interpolate(f.contiguous().numpy(), i.numpy(), w.contiguous().numpy())))


def reference_classes():
    text = open(SRC).read()
    ns = dict(torch=torch, nn=nn, F=F, List=List, pointnet2_utils=utils)
    for node in ast.parse(text).body:
        if isinstance(node, ast.ClassDef):
            exec(compile("\n".join(text.splitlines()[node.lineno - 1:node.end_lineno]), SRC, "exec"), ns)
    return ns


def randomize_bn(m):
    with torch.no_grad():
        for mod in m.modules():
            if isinstance(mod, nn.BatchNorm2d):
                mod.running_mean.normal_(0, 0.1)
                mod.running_var.uniform_(0.8, 1.2)
                mod.weight.normal_(1, 0.1)
                mod.bias.normal_(0, 0.1)


def main():
    ns = reference_classes()
    rng = np.random.default_rng(1994)
    B, N, C = 2, 700, 5
    xyz = (rng.random((B, N, 3)) * np.array([4.0, 4.0, 1.5])).astype(np.float32)
    feats = rng.standard_normal((B, C, N)).astype(np.float32)
    out = dict(xyz=xyz, feats=feats)
    torch.manual_seed(3)
    sa = ns["PointnetSAModuleMSG"](npoint=96, radii=[0.3, 0.6], nsamples=[8, 16], mlps=[[C, 16, 24], [C, 16, 32]], use_xyz=True)
    randomize_bn(sa)
    sa_all = ns["PointnetSAModule"](mlp=[C, 32], use_xyz=True, pool_method="avg_pool")     # npoint None -> GroupAll
    randomize_bn(sa_all)
    fp = ns["PointnetFPModule"](mlp=[56 + C, 32, 16])
    randomize_bn(fp)
    for name, m in (("sa", sa), ("sa_all", sa_all), ("fp", fp)):
        for k, v in m.state_dict().items():
            out["%s.%s" % (name, k)] = v.numpy().copy()
    with torch.no_grad():
        for mode in ("eval", "train"):
            for m in (sa, sa_all, fp):
                m.train(mode == "train")
            new_xyz, f1 = sa(T(xyz), T(feats))
            _, f2 = sa_all(T(xyz), T(feats))
            up = fp(T(xyz), new_xyz, T(feats), f1)
            out.update({mode + ".new_xyz": new_xyz.numpy(), mode + ".sa": f1.numpy(), mode + ".sa_all": f2.numpy(),
                        mode + ".fp": up.numpy()})
    np.savez_compressed(os.path.join(HERE, "ref_batch_modules.npz"), **out)
    print("wrote ref_batch_modules.npz", {k: v.shape for k, v in out.items() if k.startswith("eval")})


if __name__ == "__main__":
    main()
