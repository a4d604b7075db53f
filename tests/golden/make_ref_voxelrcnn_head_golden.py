"""Golden vectors for VoxelRCNNHead (roi_grid_pool + shared FC), produced by the REFERENCE class itself.

`class VoxelRCNNHead` is cut out of /root/reference/pcdet/models/roi_heads/voxelrcnn_head.py by its source lines and
executed unmodified on CPU tensors against: a bare nn.Module standing in for RoIHeadTemplate (its loss / target code is
never reached by the shipped forward), the reference's own common_utils functions (cut out the same way - the module
cannot be imported, it needs SharedArray), and the reference NeighborVoxelSAModuleMSG with the CPU oracle behind its
voxel query (make_ref_voxel_pool_golden.reference_class).  eval- and train-mode outputs are stored.
Run in the build container only:  python tests/golden/make_ref_voxelrcnn_head_golden.py -> tests/golden/ref_voxelrcnn_head.npz
"""
import ast
import importlib.util
import os
import sys
import types

import numpy as np
import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(HERE, ".."))
import cases  # noqa: E402

HEAD = "/root/reference/pcdet/models/roi_heads/voxelrcnn_head.py"
UTILS = "/root/reference/pcdet/utils/common_utils.py"
T = torch.from_numpy


def _cut(path, kind, names, ns):
    text = open(path).read()
    for node in ast.parse(text).body:
        if isinstance(node, kind) and node.name in names:
            exec(compile("\n".join(text.splitlines()[node.lineno - 1:node.end_lineno]), path, "exec"), ns)
    return ns


def reference_head():
    spec = importlib.util.spec_from_file_location("make_ref_voxel_pool_golden", os.path.join(HERE, "make_ref_voxel_pool_golden.py"))
    vp = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(vp)
    cu = _cut(UTILS, ast.FunctionDef, {"check_numpy_to_torch", "rotate_points_along_z", "get_voxel_centers",
                                       "scatter_point_inds", "generate_voxel2pinds"}, dict(torch=torch, np=np))

    class RoIHeadTemplate(nn.Module):
        def __init__(self, num_class, model_cfg, **kwargs):
            super().__init__()

    ns = dict(torch=torch, nn=nn, RoIHeadTemplate=RoIHeadTemplate, common_utils=types.SimpleNamespace(**cu),
              voxelpool_stack_modules=types.SimpleNamespace(NeighborVoxelSAModuleMSG=vp.reference_class()))
    return _cut(HEAD, ast.ClassDef, {"VoxelRCNNHead"}, ns)["VoxelRCNNHead"]


def batch_dict_of(sc, device=None, sparse=types.SimpleNamespace):
    B = sc["gt_boxes"].shape[0]
    to = (lambda a: T(np.ascontiguousarray(a))) if device is None else (lambda a: T(np.ascontiguousarray(a)).to(device))
    names = ("x_conv2", "x_conv3", "x_conv4")
    return dict(gt_boxes=to(sc["gt_boxes"]), batch_size=B,
                multi_scale_3d_strides={n: int(sc[n + ".stride"]) for n in names},
                multi_scale_3d_features={n: sparse(features=to(sc[n + ".features"]), indices=to(sc[n + ".indices"]),
                                                   spatial_shape=[int(v) for v in sc[n + ".spatial_shape"]], batch_size=B)
                                         for n in names})


def main():
    Ref = reference_head()
    sc = cases.voxelrcnn_scene()
    torch.manual_seed(1994)
    head = Ref(backbone_channels={"x_conv2": 8, "x_conv3": 12, "x_conv4": 16}, model_cfg=cases.voxelrcnn_head_cfg(),
               point_cloud_range=[float(v) for v in sc["pc_range"]], voxel_size=[float(v) for v in sc["voxel_size"]])
    out = {}
    with torch.no_grad():
        for mod in head.modules():
            if isinstance(mod, (nn.BatchNorm1d, nn.BatchNorm2d)):
                mod.running_mean.normal_(0, 0.1)
                mod.running_var.uniform_(0.8, 1.2)
                mod.weight.normal_(1, 0.1)
                mod.bias.normal_(0, 0.1)
        for k, v in head.state_dict().items():
            out["head." + k] = v.numpy().copy()
        head.eval()
        res = head(batch_dict_of(sc))
        out["eval.pooled_features"] = res["pooled_features"].numpy()
        out["eval.shared_feature"] = res["shared_feature"].numpy()
        head.train()
        out["train.pooled_features"] = head.roi_grid_pool(batch_dict_of(sc)).numpy()   # (dropout makes shared_feature random)
    np.savez_compressed(os.path.join(HERE, "ref_voxelrcnn_head.npz"), **out)
    pf = out["eval.pooled_features"]
    print("wrote ref_voxelrcnn_head.npz", pf.shape, out["eval.shared_feature"].shape,
          "non-zero grid points: %.2f" % float((np.abs(pf).sum(-1) > 0).mean()))


if __name__ == "__main__":
    main()
