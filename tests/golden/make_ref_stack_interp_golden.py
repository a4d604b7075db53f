"""Generates tests/golden/ref_stack_interp.npz by running the UNMODIFIED reference extension
(oracle/_ref/pointnet2_stack_cuda.so, compiled from /root/reference/pcdet/ops/pointnet2/pointnet2_stack/src) on a GPU:

    gpurun -- python tests/golden/make_ref_stack_interp_golden.py      # writes gpurun_out/ref_stack_interp.npz
    cp gpurun_out/ref_stack_interp.npz tests/golden/

The CPU suite (tests/test_oracle_pin.py) pins the C oracle of the stack three_nn / three_interpolate to these outputs;
inputs are rebuilt from the seed, only the kernel outputs are stored."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import oracle as O  # noqa: E402


def inputs(seed=78, ucnt=(700, 0, 333, 50), kcnt=(128, 5, 2, 77), C=9):
    """Ragged clouds incl. an empty unknown cloud, a cloud with two known points and duplicated known points (ties)."""
    rng = np.random.default_rng(seed)
    ucnt, kcnt = np.asarray(ucnt, np.int32), np.asarray(kcnt, np.int32)
    unknown = (rng.random((int(ucnt.sum()), 3)) * np.array([4.0, 4.0, 1.5])).astype(np.float32)
    known = (rng.random((int(kcnt.sum()), 3)) * np.array([4.0, 4.0, 1.5])).astype(np.float32)
    known[5:15] = known[40:50]
    feats = rng.standard_normal((known.shape[0], C)).astype(np.float32)
    weight = rng.random((unknown.shape[0], 3)).astype(np.float32)
    return unknown, ucnt, known, kcnt, feats, weight


def main():
    ref = O.load_ref("pointnet2_stack_cuda")
    dev = torch.device("cuda:0")
    unknown, ucnt, known, kcnt, feats, weight = inputs()
    u, uc, k, kc, f, w = (torch.from_numpy(a).to(dev) for a in (unknown, ucnt, known, kcnt, feats, weight))
    d2 = torch.zeros_like(u)
    idx = torch.zeros(u.shape, dtype=torch.int32, device=dev)
    ref.three_nn_wrapper(u, uc, k, kc, d2, idx)
    safe = idx.clamp(0, known.shape[0] - 1)          # the 2-point cloud leaves an index outside its own cloud; keep it in range
    out = torch.zeros((u.shape[0], feats.shape[1]), device=dev)
    ref.three_interpolate_wrapper(f, safe, w, out)
    res = dict(d2=d2.cpu().numpy(), nn_idx=idx.cpu().numpy(), interp=out.cpu().numpy())
    _, oi, od2 = O.stack_three_nn(unknown, ucnt, known, kcnt)
    print("oracle == reference:", np.array_equal(res["nn_idx"], oi) and np.array_equal(res["d2"], od2),
          np.array_equal(res["interp"], O.stack_three_interpolate(feats, np.clip(oi, 0, known.shape[0] - 1), weight)))
    os.makedirs("gpurun_out", exist_ok=True)
    np.savez_compressed("gpurun_out/ref_stack_interp.npz", **res)
    print("wrote gpurun_out/ref_stack_interp.npz")


if __name__ == "__main__":
    main()
