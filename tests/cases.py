"""Deterministic small inputs shared by the golden generator and the parity tests."""
import numpy as np

from multimodal_gar_b200 import synthetic


def frame_case(seed=1994, B=2, N=4096, M=12):
    pts, boxes = synthetic.make_batch(seed, B, N=N, M=M, density_scale=0.5, empty_frac=0.15)
    # a few adversarial boxes: zero-padded (dataloader.py:245-247), axis-aligned with a point on its face
    boxes[0, M - 1] = 0
    boxes[1, M - 2] = [1.0, 2.0, 0.0, 1.0, 1.0, 2.0, 0.0]
    pts[1, 0, :3] = [1.5, 2.0, 0.0]       # exactly on the +x face (inside: |lx| < dx/2 + margin)
    pts[1, 1, :3] = [1.0, 2.0, 1.0]       # exactly on the +z face (inside: z test is inclusive)
    pts[1, 2, :3] = [1.50002, 2.0, 0.0]   # outside by more than the 1e-5 margin
    return np.ascontiguousarray(pts[..., :3]), np.ascontiguousarray(pts[..., 3:4]), boxes


def cloud_with_ties(seed, b, n):
    """Clouds with exact duplicates (what roipoint pooling's wrap produces): FPS tie order matters."""
    rng = np.random.default_rng(seed)
    base = rng.uniform(-1, 1, (b, max(1, n // 3), 3)).astype(np.float32)
    reps = int(np.ceil(n / base.shape[1]))
    return np.ascontiguousarray(np.tile(base, (1, reps, 1))[:, :n])


def ragged_case(seed=7):
    rng = np.random.default_rng(seed)
    xyz_cnt = np.array([300, 1100, 77, 0, 513], dtype=np.int32)
    new_cnt = np.array([32, 70, 16, 0, 65], dtype=np.int32)
    xyz = rng.uniform(-1, 1, (int(xyz_cnt.sum()), 3)).astype(np.float32)
    feats = rng.uniform(0, 1, (int(xyz_cnt.sum()), 2)).astype(np.float32)
    new_xyz = rng.uniform(-1.2, 1.2, (int(new_cnt.sum()), 3)).astype(np.float32)
    return xyz, xyz_cnt, new_xyz, new_cnt, feats
