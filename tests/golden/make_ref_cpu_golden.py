"""Generates tests/golden/ref_cpu.npz from the UNMODIFIED reference points_in_boxes_cpu
(/root/reference/pcdet/ops/roiaware_pool3d/src/roiaware_pool3d.cpp:143-168, compiled by
oracle/build_ref.py into oracle/_ref).  Runs in the build container (no GPU needed):

    python oracle/build_ref.py roiaware_pool3d_cuda && python tests/golden/make_ref_cpu_golden.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from multimodal_gar_b200 import synthetic  # noqa: E402
from oracle import oracle as O  # noqa: E402


def main():
    mod = O.load_ref("roiaware_pool3d_cuda")
    pts, boxes = synthetic.make_frame(1994, N=8192, M=24)
    xyz = np.ascontiguousarray(pts[:, :3])
    xyz[0] = [boxes[3, 0], boxes[3, 1], boxes[3, 2] + boxes[3, 5] / 2]   # on the +z face of box 3
    mask = torch.zeros((boxes.shape[0], xyz.shape[0]), dtype=torch.int32)
    mod.points_in_boxes_cpu(torch.from_numpy(boxes), torch.from_numpy(xyz), mask)
    m = mask.numpy()
    out = os.path.join(ROOT, "tests", "golden", "ref_cpu.npz")
    np.savez_compressed(out, xyz=xyz, boxes=boxes, mask_bits=np.packbits(m.astype(np.uint8), axis=1),
                        hits=np.int64(m.sum()))
    print("wrote", out, os.path.getsize(out), "bytes;", int(m.sum()), "in-box pairs")


if __name__ == "__main__":
    main()
