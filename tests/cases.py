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


def voxel_scene(seed=5, B=2, grid=(6, 24, 24), voxels=1500, queries_per_frame=90, C=16, voxel_size=(0.1, 0.1, 0.2)):
    """Inputs of the voxel query path (voxel_query_utils.py / voxel_pool_modules.py): `voxels` occupied cells per
    frame of a (Z,Y,X) grid clustered in one half of it (so some queries find nothing), one point per cell,
    dense voxel -> point map, queries on and just outside the grid borders.  voxel_size is (x, y, z) metres."""
    rng = np.random.default_rng(seed)
    Z, Y, X = grid
    vs = np.asarray(voxel_size, dtype=np.float32)
    coords, cnt = [], []
    for b in range(B):
        c = np.stack([rng.integers(0, Z, voxels), rng.integers(0, Y, voxels), rng.integers(0, (2 * X) // 3, voxels)], axis=1)
        c = np.unique(c, axis=0)
        c = c[rng.permutation(len(c))]
        coords.append(np.concatenate([np.full((len(c), 1), b), c], axis=1))
        cnt.append(len(c))
    coords = np.concatenate(coords).astype(np.int32)                      # (V,4) [batch, z, y, x]
    V = len(coords)
    xyz = ((coords[:, [3, 2, 1]] + rng.random((V, 3))) * vs).astype(np.float32)
    v2p = np.full((B, Z, Y, X), -1, dtype=np.int32)
    v2p[coords[:, 0], coords[:, 1], coords[:, 2], coords[:, 3]] = np.arange(V, dtype=np.int32)
    M = queries_per_frame
    qc = np.stack([rng.integers(0, Z, B * M), rng.integers(0, Y, B * M), rng.integers(0, X, B * M)], axis=1)
    qc[:5] = 0                                                            # grid corner: clipped neighbourhood
    qc[5:10] = np.array([Z - 1, Y - 1, X - 1])                            # far corner, in the empty third
    new_coords = np.concatenate([np.repeat(np.arange(B), M)[:, None], qc], axis=1).astype(np.int32)
    new_xyz = ((qc[:, [2, 1, 0]] + rng.random((B * M, 3))) * vs).astype(np.float32)
    return dict(xyz=xyz, xyz_batch_cnt=np.asarray(cnt, dtype=np.int32), coords=coords, v2p=v2p, new_xyz=new_xyz,
                new_xyz_batch_cnt=np.full(B, M, dtype=np.int32), new_coords=new_coords,
                new_coords_xyz=np.ascontiguousarray(new_coords[:, [0, 3, 2, 1]]),
                features=rng.standard_normal((V, C)).astype(np.float32))
