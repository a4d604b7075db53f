"""Golden vectors for StackPointnetFPModule, produced by the REFERENCE class.

The class is cut out of /root/reference/pcdet/ops/pointnet2/pointnet2_stack/pointnet2_modules.py:115-157 by its source
lines and executed unmodified on CPU tensors; the two CUDA functions it calls through `pointnet2_utils` (three_nn,
three_interpolate - no GPU in the build container) are stand-ins that return the CPU oracle's results, which
tests/test_stack_interp_gpu.py pins to the reference kernels on the B200.
Run in the build container only:  python tests/golden/make_ref_stack_fp_golden.py -> tests/golden/ref_stack_fp.npz
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

SRC = "/root/reference/pcdet/ops/pointnet2/pointnet2_stack/pointnet2_modules.py"
T = torch.from_numpy


def _three_nn(unknown, ucnt, known, kcnt):
    d, i, _ = O.stack_three_nn(unknown.numpy(), ucnt.numpy(), known.numpy(), kcnt.numpy())
    return T(d), T(i)


utils = types.SimpleNamespace(
    three_nn=_three_nn,
    three_interpolate=lambda f, i, w: T(O.stack_three_interpolate(f.contiguous().numpy(), i.numpy(), w.contiguous().numpy())))


def reference_class(name):
    text = open(SRC).read()
    ns = dict(torch=torch, nn=nn, F=F, List=List, pointnet2_utils=utils)
    for node in ast.parse(text).body:
        if isinstance(node, ast.ClassDef) and node.name == name:
            exec(compile("\n".join(text.splitlines()[node.lineno - 1:node.end_lineno]), SRC, "exec"), ns)
    return ns[name]


def inputs(seed=1994, ucnt=(500, 301, 64), kcnt=(96, 40, 3), C1=5, C2=24):
    rng = np.random.default_rng(seed)
    ucnt, kcnt = np.asarray(ucnt, np.int32), np.asarray(kcnt, np.int32)
    unknown = (rng.random((int(ucnt.sum()), 3)) * np.array([4.0, 4.0, 1.5])).astype(np.float32)
    known = (rng.random((int(kcnt.sum()), 3)) * np.array([4.0, 4.0, 1.5])).astype(np.float32)
    known[:7] = unknown[:7]                                     # coincident points: zero distances, 1e-8 regulariser
    ufeat = rng.standard_normal((unknown.shape[0], C1)).astype(np.float32)
    kfeat = rng.standard_normal((known.shape[0], C2)).astype(np.float32)
    return unknown, ucnt, known, kcnt, ufeat, kfeat


def main():
    cls = reference_class("StackPointnetFPModule")
    unknown, ucnt, known, kcnt, ufeat, kfeat = inputs()
    torch.manual_seed(5)
    fp = cls(mlp=[kfeat.shape[1] + ufeat.shape[1], 32, 16])
    fp_nofeat = cls(mlp=[kfeat.shape[1], 8])
    out = {}
    with torch.no_grad():
        for m in (fp, fp_nofeat):
            for mod in m.modules():
                if isinstance(mod, nn.BatchNorm2d):
                    mod.running_mean.normal_(0, 0.1)
                    mod.running_var.uniform_(0.8, 1.2)
                    mod.weight.normal_(1, 0.1)
                    mod.bias.normal_(0, 0.1)
        for name, m in (("fp", fp), ("fp_nofeat", fp_nofeat)):
            for k, v in m.state_dict().items():
                out["%s.%s" % (name, k)] = v.numpy().copy()
        for mode in ("eval", "train"):
            fp.train(mode == "train")
            fp_nofeat.train(mode == "train")
            out[mode + ".fp"] = fp(T(unknown), T(ucnt), T(known), T(kcnt), T(ufeat), T(kfeat)).numpy()
            out[mode + ".fp_nofeat"] = fp_nofeat(T(unknown), T(ucnt), T(known), T(kcnt), None, T(kfeat)).numpy()
    np.savez_compressed(os.path.join(HERE, "ref_stack_fp.npz"), **out)
    print("wrote ref_stack_fp.npz", {k: v.shape for k, v in out.items() if k.startswith("eval")})


if __name__ == "__main__":
    main()
