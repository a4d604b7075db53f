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


class AttrDict(dict):
    """dict with attribute access (what the reference gets from EasyDict for its yaml configs)."""
    __getattr__ = dict.__getitem__

    @staticmethod
    def of(obj):
        if isinstance(obj, dict):
            return AttrDict({k: AttrDict.of(v) for k, v in obj.items()})
        return [AttrDict.of(v) for v in obj] if isinstance(obj, list) else obj


def voxelrcnn_head_cfg(channels=16, grid_size=4):
    """ROI_HEAD block of Multimodal_cfg/mil3.yaml:105-135 at test size (same structure, smaller widths)."""
    layers = {name: dict(MLPS=[[channels, channels]], QUERY_RANGES=[[4, 4, 4]], POOL_RADIUS=[r], NSAMPLE=[16],
                         POOL_METHOD="max_pool")
              for name, r in (("x_conv2", 0.4), ("x_conv3", 0.8), ("x_conv4", 1.6))}
    return AttrDict.of(dict(NAME="VoxelRCNNHead", CLASS_AGNOSTIC=True, SHARED_FC=[64, 32], DP_RATIO=0.3,
                            ROI_GRID_POOL=dict(FEATURES_SOURCE=["x_conv2", "x_conv3", "x_conv4"], PRE_MLP=True,
                                               GRID_SIZE=grid_size, POOL_LAYERS=layers)))


def voxelrcnn_scene(seed=11, B=2, rois=4, grid=(16, 96, 96), voxel_size=(0.1, 0.1, 0.2), channels=(8, 12, 16)):
    """Inputs of VoxelRCNNHead.roi_grid_pool: person boxes (one reaching outside the point-cloud range, one far from
    every voxel) and three scales (strides 2, 4, 8) of occupied voxels with features, clustered around the boxes."""
    rng = np.random.default_rng(seed)
    Z, Y, X = grid
    vs = np.asarray(voxel_size, dtype=np.float32)
    pc_range = [0.0, 0.0, 0.0, float(X * vs[0]), float(Y * vs[1]), float(Z * vs[2])]
    boxes = np.zeros((B, rois, 8), dtype=np.float32)
    boxes[..., 0] = rng.uniform(1.5, 8.0, (B, rois))
    boxes[..., 1] = rng.uniform(1.5, 8.0, (B, rois))
    boxes[..., 2] = rng.uniform(1.0, 2.0, (B, rois))
    boxes[..., 3:5] = rng.uniform(0.4, 0.9, (B, rois, 2))
    boxes[..., 5] = rng.uniform(1.4, 1.95, (B, rois))
    boxes[..., 6] = rng.uniform(-np.pi, np.pi, (B, rois))
    boxes[..., 7] = 1.0                                                   # class column: rois are (…, 7 + C)
    boxes[0, 0, :2] = (0.1, 0.2)                                          # grid points with negative voxel coordinates
    boxes[1, 1, :2] = (9.3, 9.4)                                          # nothing nearby
    out = dict(gt_boxes=boxes, pc_range=np.asarray(pc_range, np.float32), voxel_size=vs)
    for name, stride, C in zip(("x_conv2", "x_conv3", "x_conv4"), (2, 4, 8), channels):
        shape = (Z // stride, Y // stride, X // stride)
        coords = []
        for b in range(B):
            ctr = boxes[b, rng.integers(0, rois, 4000 // stride), :3] + rng.normal(0, 0.5, (4000 // stride, 3))
            c = np.floor(ctr / (vs * stride)).astype(np.int64)[:, ::-1]    # -> (z, y, x)
            c = c[((c >= 0) & (c < np.asarray(shape))).all(axis=1)]
            c = np.unique(c, axis=0)
            c = c[rng.permutation(len(c))]
            coords.append(np.concatenate([np.full((len(c), 1), b), c], axis=1))
        coords = np.concatenate(coords).astype(np.int32)
        out[name + ".indices"] = coords
        out[name + ".features"] = rng.standard_normal((len(coords), C)).astype(np.float32)
        out[name + ".spatial_shape"] = np.asarray(shape, np.int32)
        out[name + ".stride"] = np.asarray(stride, np.int32)
    return out


def unambiguous(O, xyz, boxes, max_amb=8):
    """(point_ok (N,), box_ok (M,)) for one frame: the CPU oracle uses glibc cosf/sinf, the GPU libdevice, so a
    (point, box) pair whose in-box decision flips under a 1-ulp trig change may legitimately differ.  Such pairs are
    rare (a point within ~1e-6 m of a face): comparisons keep every point / box that has none instead of dropping the
    whole frame, and the count is asserted small."""
    amb = O.in_box_ambiguous(np.ascontiguousarray(xyz), np.ascontiguousarray(boxes)) != 0
    assert int(amb.sum()) <= max_amb, "suspiciously many trig-ambiguous pairs: %d" % int(amb.sum())
    return ~amb.any(axis=0), ~amb.any(axis=1)


def canonical_close(got, want, pooled_raw, boxes, rel=1e-5):
    """Canonical-frame coordinates (pointrcnn_head.py:121-129): |got - want| <= rel * |(x - cx, y - cy)| per point -
    north_star's 1e-5 relative, taken against the rotated VECTOR (a rotation cancels, so an element can be ~0 while its
    error is an ulp of the vector's length); z is a plain subtraction and the features are copies: exact.
    got / want (M,S,3+C) canonical, pooled_raw (M,S,3+C) sensor frame, boxes (M,7)."""
    sx = pooled_raw[..., 0] - boxes[:, None, 0]
    sy = pooled_raw[..., 1] - boxes[:, None, 1]
    bound = rel * np.hypot(sx, sy) + 1e-30
    err = np.abs(got[..., :2].astype(np.float64) - want[..., :2].astype(np.float64)).max(axis=-1)
    assert (err <= bound).all(), "canonical xy off by %g of the vector length" % float((err / bound).max() * rel)
    np.testing.assert_array_equal(got[..., 2:], want[..., 2:])
