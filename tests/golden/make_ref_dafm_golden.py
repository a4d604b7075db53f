"""Golden vectors for the Distance Aware Fusion Module, produced by the REFERENCE class itself.

`FusionAttention_mat` is cut out of /root/reference/model/gat_model.py by its source lines (the module cannot be
imported: it needs torchmetrics / torch_geometric / spconv, absent here) and executed unmodified on seeded
inputs; Dg / De come from torchvision.ops.generalized_box_iou (what the reference calls, gat_model.py:1468) and
from the definition of the pairwise euclidean distance (torchmetrics is unpinned and absent).
Run in the build container only:   python tests/golden/make_ref_dafm_golden.py   -> tests/golden/ref_dafm.npz
"""
import ast
import math
import os

import numpy as np
import torch
import torch.nn as nn
import torchvision.ops as TO

SRC = "/root/reference/model/gat_model.py"
HERE = os.path.dirname(os.path.abspath(__file__))


def reference_class():
    text = open(SRC).read()
    node = [n for n in ast.parse(text).body if isinstance(n, ast.ClassDef) and n.name == "FusionAttention_mat"][0]
    code = "\n".join(text.splitlines()[node.lineno - 1:node.end_lineno])
    ns = dict(torch=torch, nn=nn, math=math)
    exec(compile(code, SRC, "exec"), ns)
    return ns["FusionAttention_mat"]


def main():
    torch.manual_seed(1994)
    Ref = reference_class()
    D, counts, MNP = 64, [7, 12, 1, 12], 12
    m1, m2 = Ref(D, D), Ref(D, D)
    with torch.no_grad():
        for m in (m1, m2):
            for p in m.parameters():
                if p.dim() == 1:
                    p.normal_(0.5, 0.3)
    out = dict(counts=np.array(counts), D=D)
    for k, v in list(m1.state_dict().items()):
        out["m1." + k] = v.numpy()
    for k, v in list(m2.state_dict().items()):
        out["m2." + k] = v.numpy()
    B = len(counts)
    rgb, lidar = torch.randn(B, MNP, D), torch.randn(B, MNP, D)
    xy = torch.rand(B, MNP, 2) * 300
    wh = torch.rand(B, MNP, 2) * 80 + 5
    bboxes = torch.cat([xy, xy + wh], dim=-1)
    bboxes3d = torch.cat([torch.randn(B, MNP, 3) * 6, torch.rand(B, MNP, 4)], dim=-1)
    fused = torch.zeros(B, MNP, D)
    with torch.no_grad():
        for b, n in enumerate(counts):                       # gat_model.py:1396-1405, :1467-1475
            Rb, Lb = rgb[b, :n], lidar[b, :n]
            Dg = TO.generalized_box_iou(bboxes[b, :n], bboxes[b, :n])
            c = bboxes3d[b, :n, :3].double()
            De = torch.cdist(c, c).float()
            De.fill_diagonal_(0)
            Rp, Lp = m1(Rb, Lb, Dg, De)
            Rp, Lp = m2(Rp, Lp, Dg, De)
            fused[b, :n], _ = torch.max(torch.stack((Rp, Lp)), dim=0)
    out.update(rgb=rgb.numpy(), lidar=lidar.numpy(), bboxes=bboxes.numpy(), bboxes3d=bboxes3d.numpy(), fused=fused.numpy())
    np.savez_compressed(os.path.join(HERE, "ref_dafm.npz"), **out)
    print("wrote ref_dafm.npz", {k: getattr(v, "shape", v) for k, v in out.items() if not k.startswith("m")})


if __name__ == "__main__":
    main()
