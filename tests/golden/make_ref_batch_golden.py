"""Generates tests/golden/ref_batch_ops.npz by running the UNMODIFIED reference extension
(oracle/_ref/pointnet2_batch_cuda.so, compiled from /root/reference/pcdet/ops/pointnet2/pointnet2_batch/src) on a GPU:

    gpurun -- python tests/golden/make_ref_batch_golden.py      # writes gpurun_out/ref_batch_ops.npz
    cp gpurun_out/ref_batch_ops.npz tests/golden/

The CPU suite (tests/test_oracle_pin.py) pins the C oracle of the batch ball query, three_nn and three_interpolate to
these outputs; inputs are rebuilt from the seed, only the kernel outputs are stored."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import oracle as O  # noqa: E402


def inputs(seed=77, B=2, N=900, M=128, C=7):
    rng = np.random.default_rng(seed)
    xyz = (rng.random((B, N, 3)) * np.array([4.0, 4.0, 1.5])).astype(np.float32)
    new_xyz = np.ascontiguousarray(xyz[:, rng.permutation(N)[:M]])
    new_xyz[:, :2] += 50.0
    feats = rng.standard_normal((B, C, M)).astype(np.float32)
    weight = rng.random((B, N, 3)).astype(np.float32)
    return xyz, new_xyz, feats, weight


def main():
    ref = O.load_ref("pointnet2_batch_cuda")
    dev = torch.device("cuda:0")
    xyz, new_xyz, feats, weight = inputs()
    B, N, M, C = xyz.shape[0], xyz.shape[1], new_xyz.shape[1], feats.shape[1]
    x, q, f, w = (torch.from_numpy(a).to(dev) for a in (xyz, new_xyz, feats, weight))
    bq = torch.zeros((B, M, 16), dtype=torch.int32, device=dev)
    ref.ball_query_wrapper(B, N, M, 0.35, 16, q, x, bq)
    d2 = torch.zeros((B, N, 3), device=dev)
    nn_idx = torch.zeros((B, N, 3), dtype=torch.int32, device=dev)
    ref.three_nn_wrapper(B, N, M, x, q, d2, nn_idx)
    out = torch.zeros((B, C, N), device=dev)
    ref.three_interpolate_wrapper(B, C, M, N, f, nn_idx, w, out)
    res = dict(bq=bq.cpu().numpy(), d2=d2.cpu().numpy(), nn_idx=nn_idx.cpu().numpy(), interp=out.cpu().numpy())
    _, oi, od2 = O.three_nn(xyz, new_xyz)
    print("oracle == reference:", np.array_equal(res["bq"], O.batch_ball_query(0.35, 16, xyz, new_xyz)),
          np.array_equal(res["nn_idx"], oi) and np.array_equal(res["d2"], od2),
          np.array_equal(res["interp"], O.three_interpolate(feats, oi, weight)))
    os.makedirs("gpurun_out", exist_ok=True)
    np.savez_compressed("gpurun_out/ref_batch_ops.npz", **res)
    print("wrote gpurun_out/ref_batch_ops.npz")


if __name__ == "__main__":
    main()
