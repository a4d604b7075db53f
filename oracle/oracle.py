"""TEST INFRASTRUCTURE ONLY - Python face of the CPU oracle.

numpy/ctypes wrappers over oracle/_build/libmgar_oracle.so (plain-C restatement, see
mgar_oracle.c), torch-CPU restatements of the pure-torch parts of the reference path
(QueryAndGroup / StackSAModuleMSG / canonical transform), and a loader for the UNMODIFIED
reference extensions compiled into oracle/_ref/ (oracle/build_ref.py).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  Citations are relative to /root/reference/.
"""
import ctypes
import importlib.util
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

_f = ctypes.POINTER(ctypes.c_float)
_i = ctypes.POINTER(ctypes.c_int)


def build():
    """Compile the C restatement (gcc is the only requirement)."""
    subprocess.check_call(["make", "-s", "-C", HERE])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(HERE, "_build", "libmgar_oracle.so")
        src = os.path.join(HERE, "mgar_oracle.c")
        if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(src):
            build()
        _LIB = ctypes.CDLL(path)
    return _LIB


def _fp(a):
    assert a.dtype == np.float32 and a.flags.c_contiguous
    return a.ctypes.data_as(_f)


def _ip(a):
    assert a.dtype == np.int32 and a.flags.c_contiguous
    return a.ctypes.data_as(_i)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


# ---- in-box ------------------------------------------------------------------------------
def points_in_boxes_gpu(points, boxes):
    """(B,N,3),(B,M,7) -> (B,N) int32 first-hit box or -1. pcdet/ops/roiaware_pool3d/roiaware_pool3d_utils.py:28-41"""
    points, boxes = _f32(points), _f32(boxes)
    B, N, _ = points.shape
    M = boxes.shape[1]
    out = np.full((B, N), -1, dtype=np.int32)
    lib().orc_points_in_boxes_first(_fp(boxes), _fp(points), _ip(out), B, M, N)
    return out


def points_in_boxes_cpu(points, boxes):
    """(N,3),(M,7) -> (M,N) int32 0/1, margin 1e-2. roiaware_pool3d_utils.py:9-25"""
    points, boxes = _f32(points), _f32(boxes)
    out = np.zeros((boxes.shape[0], points.shape[0]), dtype=np.int32)
    lib().orc_points_in_boxes_mask(_fp(boxes), _fp(points), _ip(out), boxes.shape[0], points.shape[0])
    return out


def in_box_mask_gpu_sem(points, boxes):
    points, boxes = _f32(points), _f32(boxes)
    out = np.zeros((boxes.shape[0], points.shape[0]), dtype=np.int32)
    lib().orc_in_box_mask_gpu_sem(_fp(boxes), _fp(points), _ip(out), boxes.shape[0], points.shape[0])
    return out


def in_box_ambiguous(points, boxes, cpu_sem=False):
    """(M,N) 0/1: pairs whose decision flips under a 1-ulp change of cosf/sinf."""
    points, boxes = _f32(points), _f32(boxes)
    out = np.zeros((boxes.shape[0], points.shape[0]), dtype=np.int32)
    lib().orc_in_box_ambiguous(_fp(boxes), _fp(points), _ip(out), boxes.shape[0], points.shape[0], int(cpu_sem))
    return out


def enlarge_box3d(boxes, extra_width):
    """pcdet/utils/box_utils.py:187-201 (dx,dy,dz += extra, fp32)."""
    boxes = _f32(boxes).copy()
    boxes[..., 3:6] = boxes[..., 3:6] + np.asarray(extra_width, dtype=np.float32)
    return boxes


def roipoint_pool3d(points, point_features, boxes3d, pool_extra_width, num_sampled_points=512, return_idx=False):
    """pcdet/ops/roipoint_pool3d/roipoint_pool3d_utils.py:33-59 -> pooled (B,M,S,3+C), empty (B,M)."""
    points, feats = _f32(points), _f32(point_features)
    B, N, _ = points.shape
    M, C, S = boxes3d.shape[1], feats.shape[2], num_sampled_points
    boxes = enlarge_box3d(boxes3d, pool_extra_width)
    pooled = np.zeros((B, M, S, 3 + C), dtype=np.float32)
    empty = np.zeros((B, M), dtype=np.int32)
    pidx = np.full((B, M, S), -1, dtype=np.int32)
    lib().orc_roipoint_pool(_fp(points), _fp(boxes), _fp(feats), _fp(pooled), _ip(empty), _ip(pidx), B, N, M, S, C)
    return (pooled, empty, pidx) if return_idx else (pooled, empty)


# ---- FPS ---------------------------------------------------------------------------------
def roiaware_pool3d(rois, pts, pts_feature, out_size, max_pts_each_voxel=128, pool_method="max"):
    """RoIAwarePool3dFunction.forward, pcdet/ops/roiaware_pool3d/roiaware_pool3d_utils.py:54-92
    -> pooled (N,ox,oy,oz,C), argmax (same, int32; zeros for avg), pts_idx_of_voxels (N,ox,oy,oz,max_pts)."""
    rois, pts, feat = _f32(rois), _f32(pts), _f32(pts_feature)
    ox, oy, oz = (out_size,) * 3 if isinstance(out_size, int) else out_size
    Nb, P, C = rois.shape[0], pts.shape[0], feat.shape[1]
    pooled = np.zeros((Nb, ox, oy, oz, C), dtype=np.float32)
    argmax = np.zeros((Nb, ox, oy, oz, C), dtype=np.int32)
    lists = np.zeros((Nb, ox, oy, oz, max_pts_each_voxel), dtype=np.int32)
    lib().orc_roiaware_pool3d(_fp(rois), _fp(pts), _fp(feat), _ip(argmax), _ip(lists), _fp(pooled), Nb, P, C,
                              max_pts_each_voxel, ox, oy, oz, {"max": 0, "avg": 1}[pool_method])
    return pooled, argmax, lists


def roiaware_pool3d_backward(lists, argmax, grad_out, num_pts, pool_method="max"):
    """RoIAwarePool3dFunction.backward, roiaware_pool3d_utils.py:94-107 -> grad_in (num_pts, C)."""
    lists, argmax, grad_out = _i32(lists), _i32(argmax), _f32(grad_out)
    Nb, ox, oy, oz, max_pts = lists.shape
    C = grad_out.shape[-1]
    grad_in = np.zeros((num_pts, C), dtype=np.float32)
    lib().orc_roiaware_pool3d_backward(_ip(lists), _ip(argmax), _fp(grad_out), _fp(grad_in), Nb, ox, oy, oz, C, max_pts,
                                       {"max": 0, "avg": 1}[pool_method])
    return grad_in


def opt_n_threads(n):
    return lib().orc_opt_n_threads(int(n))


def farthest_point_sample(xyz, npoint, temp=None, return_temp=False):
    """pointnet2_stack/pointnet2_utils.py:162-188: xyz (B,N,3) -> idx (B,npoint) int32."""
    xyz = _f32(xyz)
    B, N, _ = xyz.shape
    temp = np.full((B, N), 1e10, dtype=np.float32) if temp is None else _f32(temp).copy()
    idx = np.zeros((B, npoint), dtype=np.int32)
    lib().orc_fps_batch(B, N, npoint, _fp(xyz), _fp(temp), _ip(idx))
    return (idx, temp) if return_temp else idx


def stack_farthest_point_sample(xyz, xyz_batch_cnt, npoint, return_temp=False):
    """pointnet2_utils.py:191-225: xyz (sumN,3) -> idx (sum npoint) GLOBAL indices."""
    xyz, cnt = _f32(xyz), _i32(xyz_batch_cnt)
    if np.isscalar(npoint):
        npoint = [npoint] * len(cnt)
    npoint = _i32(npoint)
    temp = np.full((xyz.shape[0],), 1e10, dtype=np.float32)
    idx = np.zeros((int(npoint.sum()),), dtype=np.int32)
    lib().orc_fps_stack(len(cnt), _fp(xyz), _fp(temp), _ip(cnt), _ip(idx), _ip(npoint))
    return (idx, temp) if return_temp else idx


# ---- ball query / grouping -----------------------------------------------------------------
def ball_query_raw(radius, nsample, xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt):
    """Raw kernel output (with the -1 sentinel), idx zero-filled first. ball_query_gpu.cu:16-66"""
    xyz, new_xyz = _f32(xyz), _f32(new_xyz)
    cnt, ncnt = _i32(xyz_batch_cnt), _i32(new_xyz_batch_cnt)
    M = new_xyz.shape[0]
    idx = np.zeros((M, nsample), dtype=np.int32)
    lib().orc_ball_query_stack(len(cnt), M, ctypes.c_float(radius), nsample, _fp(new_xyz), _ip(ncnt),
                               _fp(xyz), _ip(cnt), _ip(idx))
    return idx


def ball_query(radius, nsample, xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt):
    """pointnet2_utils.py:8-42 -> (idx with empty rows zeroed, empty_ball_mask)."""
    idx = ball_query_raw(radius, nsample, xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt)
    mask = idx[:, 0] == -1
    idx[mask] = 0
    return idx, mask


def voxel_query_raw(max_range, radius, nsample, xyz, new_xyz, new_coords, point_indices):
    """pointnet2_stack_cuda.voxel_query_wrapper as VoxelQuery.forward calls it
    (pcdet/ops/pointnet2/pointnet2_stack/voxel_query_utils.py:30-37): raw (M,nsample) rows, global indices."""
    xyz, new_xyz = _f32(xyz), _f32(new_xyz)
    new_coords, point_indices = _i32(new_coords), _i32(point_indices)
    M = new_coords.shape[0]
    _, Z, Y, X = point_indices.shape
    idx = np.zeros((M, nsample), dtype=np.int32)
    lib().orc_voxel_query_stack(M, Z, Y, X, nsample, ctypes.c_float(radius), max_range[0], max_range[1], max_range[2],
                                _fp(new_xyz), _fp(xyz), _ip(new_coords), _ip(point_indices), _ip(idx))
    return idx


def voxel_query(max_range, radius, nsample, xyz, new_xyz, new_coords, point_indices):
    """VoxelQuery.forward, voxel_query_utils.py:12-44 -> idx (rows without neighbours zeroed), empty mask."""
    idx = voxel_query_raw(max_range, radius, nsample, xyz, new_xyz, new_coords, point_indices)
    empty = idx[:, 0] == -1
    idx[empty] = 0
    return idx, empty


def voxel_query_and_group(max_range, radius, nsample, new_coords, xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt,
                          features, point_indices):
    """VoxelQueryAndGrouping.forward, voxel_query_utils.py:63-100 -> grouped_features (M,C,ns), grouped_xyz (M,3,ns),
    empty (M)."""
    idx, empty = voxel_query(max_range, radius, nsample, xyz, new_xyz, new_coords, point_indices)
    B = len(xyz_batch_cnt)
    idx = idx.reshape(B, -1, nsample)
    count = 0
    for b in range(B):
        idx[b] -= count
        count += int(xyz_batch_cnt[b])
    idx = idx.reshape(-1, nsample)
    idx[empty] = 0
    return (grouping_operation(features, xyz_batch_cnt, idx, new_xyz_batch_cnt),
            grouping_operation(xyz, xyz_batch_cnt, idx, new_xyz_batch_cnt), empty)


def grouping_operation(features, features_batch_cnt, idx, idx_batch_cnt):
    """pointnet2_utils.py:52-86 -> (M,C,nsample)."""
    features, idx = _f32(features), _i32(idx)
    fcnt, icnt = _i32(features_batch_cnt), _i32(idx_batch_cnt)
    M, ns = idx.shape
    C = features.shape[1]
    out = np.empty((M, C, ns), dtype=np.float32)
    lib().orc_group_points_stack(len(icnt), M, C, ns, _fp(features), _ip(fcnt), _ip(idx), _ip(icnt), _fp(out))
    return out


def grouping_operation_grad(grad_out, idx, idx_batch_cnt, features_batch_cnt, N):
    """pointnet2_utils.py:88-106 -> (N,C)."""
    grad_out, idx = _f32(grad_out), _i32(idx)
    fcnt, icnt = _i32(features_batch_cnt), _i32(idx_batch_cnt)
    M, C, ns = grad_out.shape
    g = np.zeros((N, C), dtype=np.float32)
    lib().orc_group_points_grad_stack(len(icnt), M, C, N, ns, _fp(grad_out), _ip(idx), _ip(icnt), _ip(fcnt), _fp(g))
    return g


def query_and_group(radius, nsample, xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt, features=None, use_xyz=True):
    """pointnet2_utils.py:123-159 -> (new_features (M,3+C,ns), idx)."""
    idx, mask = ball_query(radius, nsample, xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt)
    gx = grouping_operation(xyz, xyz_batch_cnt, idx, new_xyz_batch_cnt)
    gx = gx - _f32(new_xyz)[:, :, None]
    gx[mask] = 0
    if features is None:
        return gx, idx
    gf = grouping_operation(features, xyz_batch_cnt, idx, new_xyz_batch_cnt)
    gf[mask] = 0
    return (np.concatenate([gx, gf], axis=1) if use_xyz else gf), idx


# ---- canonical transform (torch ops in the reference) ------------------------------------------
def canonical_transform(pooled, boxes3d, empty_flag):
    """pcdet/models/roi_heads/pointrcnn_head.py:121-129 with utils/common_utils.py:35-57.

    pooled (B,M,S,3+C) -> (B*M,S,3+C): xyz -= centre; xyz = xyz @ Rz(-heading); empty rows zeroed.
    The reference does the rotation as a batched fp32 torch.matmul whose summation order is the
    BLAS library's; this restatement fixes it as x' = x*c + y*(-s), y' = x*s + y*c in float32
    (numpy, separate roundings) - compared with a 1e-5 relative tolerance, not bit-exactly.
    """
    pooled = _f32(pooled).copy()
    boxes = _f32(boxes3d)
    B, M, S, D = pooled.shape
    pooled[..., 0:3] -= boxes[:, :, None, 0:3]
    pooled = pooled.reshape(B * M, S, D)
    ang = (-boxes.reshape(B * M, -1)[:, 6]).astype(np.float32)
    c, s = np.cos(ang).astype(np.float32)[:, None], np.sin(ang).astype(np.float32)[:, None]
    x, y = pooled[:, :, 0].copy(), pooled[:, :, 1].copy()
    pooled[:, :, 0] = x * c + y * (-s)
    pooled[:, :, 1] = x * s + y * c
    pooled[np.asarray(empty_flag).reshape(-1) > 0] = 0
    return pooled


# ---- SA shared MLP (torch-CPU fp32, the authoritative feature oracle) --------------------------
def make_sa_mlp(mlp_spec, seed=0):
    """The reference's shared-MLP container: pointnet2_stack/pointnet2_modules.py:56-76.
    mlp_spec already includes the +3 for use_xyz.  kaiming_normal conv, BN gamma=1 beta=0; to
    make eval-mode BN non-trivial the running stats / affine are then randomised (seeded)."""
    import torch
    import torch.nn as nn
    g = torch.Generator().manual_seed(seed)
    layers = []
    for k in range(len(mlp_spec) - 1):
        conv = nn.Conv2d(mlp_spec[k], mlp_spec[k + 1], kernel_size=1, bias=False)
        bn = nn.BatchNorm2d(mlp_spec[k + 1])
        with torch.no_grad():
            std = (2.0 / mlp_spec[k]) ** 0.5
            conv.weight.copy_(torch.randn(conv.weight.shape, generator=g) * std)
            bn.weight.copy_(1.0 + 0.1 * torch.randn(bn.weight.shape, generator=g))
            bn.bias.copy_(0.1 * torch.randn(bn.bias.shape, generator=g))
            bn.running_mean.copy_(0.1 * torch.randn(bn.running_mean.shape, generator=g))
            bn.running_var.copy_(1.0 + 0.2 * torch.rand(bn.running_var.shape, generator=g))
        layers.extend([conv, bn, nn.ReLU()])
    return nn.Sequential(*layers).eval()


def sa_mlp_forward(mlp, new_features, pool_method="max_pool"):
    """pointnet2_modules.py:93-106: (M,C0,ns) -> (M,C_last), torch CPU fp32, eval or train mode."""
    import torch
    import torch.nn.functional as F
    x = torch.as_tensor(np.asarray(new_features), dtype=torch.float32)
    x = x.permute(1, 0, 2).unsqueeze(0)
    with torch.no_grad():
        y = mlp(x)
        if pool_method == "max_pool":
            y = F.max_pool2d(y, kernel_size=[1, y.size(3)]).squeeze(-1)
        else:
            y = F.avg_pool2d(y, kernel_size=[1, y.size(3)]).squeeze(-1)
    return y.squeeze(0).permute(1, 0).contiguous().numpy()


def stack_sa_module_msg(mlps, radii, nsamples, xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt, features=None):
    """StackSAModuleMSG.forward, pointnet2_modules.py:78-112 (use_xyz=True, max_pool)."""
    outs = []
    for mlp, r, ns in zip(mlps, radii, nsamples):
        nf, _ = query_and_group(r, ns, xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt, features)
        outs.append(sa_mlp_forward(mlp, nf))
    return np.concatenate(outs, axis=1)


# ---- the unmodified reference, compiled into oracle/_ref (see build_ref.py) ---------------------
# ---- pointnet2_batch surface: (B,N,3) coordinates, channel-major (B,C,N) features -------------
def batch_ball_query(radius, nsample, xyz, new_xyz):
    """BallQuery.forward, pcdet/ops/pointnet2/pointnet2_batch/pointnet2_utils.py:193-214 -> (B,npoint,nsample) int32."""
    xyz, new_xyz = _f32(xyz), _f32(new_xyz)
    B, N, _ = xyz.shape
    M = new_xyz.shape[1]
    idx = np.zeros((B, M, nsample), dtype=np.int32)
    lib().orc_ball_query_batch(B, N, M, ctypes.c_float(radius), nsample, _fp(new_xyz), _fp(xyz), _ip(idx))
    return idx


def batch_grouping_operation(features, idx):
    """GroupingOperation.forward, pointnet2_batch/pointnet2_utils.py:153-170: (B,C,N), (B,npoint,nsample) ->
    (B,C,npoint,nsample).  GatherOperation (:39-55) is the nsample-free case."""
    features, idx = _f32(features), _i32(idx)
    B, C, _ = features.shape
    flat = idx.reshape(B, 1, -1).astype(np.int64)
    return np.take_along_axis(features, np.broadcast_to(flat, (B, C, flat.shape[2])), axis=2).reshape((B, C) + idx.shape[1:])


def batch_grouping_grad(grad_out, idx, N):
    """GroupingOperation.backward / GatherOperation.backward (:172-187, :57-69): scatter-add into (B,C,N)."""
    grad_out, idx = _f32(grad_out), _i32(idx)
    B, C = grad_out.shape[:2]
    g = np.zeros((B, C, N), dtype=np.float32)
    for b in range(B):
        for c in range(C):
            np.add.at(g[b, c], idx[b].reshape(-1), grad_out[b, c].reshape(-1))
    return g


def three_nn(unknown, known):
    """ThreeNN.forward, pointnet2_batch/pointnet2_utils.py:75-96 -> dist (B,N,3) (sqrt of the kernel's output), idx."""
    unknown, known = _f32(unknown), _f32(known)
    B, N, _ = unknown.shape
    dist2 = np.zeros((B, N, 3), dtype=np.float32)
    idx = np.zeros((B, N, 3), dtype=np.int32)
    lib().orc_three_nn(B, N, known.shape[1], _fp(unknown), _fp(known), _fp(dist2), _ip(idx))
    return np.sqrt(dist2), idx, dist2


def three_interpolate(features, idx, weight):
    """ThreeInterpolate.forward, pointnet2_batch/pointnet2_utils.py:107-128: (B,C,M), (B,N,3) x2 -> (B,C,N)."""
    features, idx, weight = _f32(features), _i32(idx), _f32(weight)
    B, C, M = features.shape
    N = idx.shape[1]
    out = np.zeros((B, C, N), dtype=np.float32)
    lib().orc_three_interpolate(B, C, M, N, _fp(features), _ip(idx), _fp(weight), _fp(out))
    return out


def three_interpolate_grad(grad_out, idx, weight, M):
    """ThreeInterpolate.backward (:130-147): grad_out (B,C,N) * weight scattered into (B,C,M)."""
    grad_out, idx, weight = _f32(grad_out), _i32(idx), _f32(weight)
    B, C, N = grad_out.shape
    g = np.zeros((B, C, M), dtype=np.float32)
    for b in range(B):
        for c in range(C):
            for j in range(3):
                np.add.at(g[b, c], idx[b, :, j], grad_out[b, c] * weight[b, :, j])
    return g


def stack_three_nn(unknown, unknown_batch_cnt, known, known_batch_cnt):
    """ThreeNN.forward, pointnet2_stack/pointnet2_utils.py:228-258 -> dist (N,3) (sqrt of the kernel's output),
    idx (N,3) global known rows, and the kernel's squared distances."""
    unknown, known = _f32(unknown), _f32(known)
    ucnt, kcnt = _i32(unknown_batch_cnt), _i32(known_batch_cnt)
    assert ucnt.sum() == unknown.shape[0] and kcnt.sum() == known.shape[0] and len(ucnt) == len(kcnt)
    dist2 = np.zeros(unknown.shape, dtype=np.float32)
    idx = np.zeros(unknown.shape, dtype=np.int32)
    lib().orc_three_nn_stack(len(ucnt), _fp(unknown), _ip(ucnt), _fp(known), _ip(kcnt), _fp(dist2), _ip(idx))
    return np.sqrt(dist2), idx, dist2


def stack_three_interpolate(features, idx, weight):
    """ThreeInterpolate.forward, pointnet2_stack/pointnet2_utils.py:264-284: (M,C), (N,3) x2 -> (N,C)."""
    features, idx, weight = _f32(features), _i32(idx), _f32(weight)
    out = np.zeros((idx.shape[0], features.shape[1]), dtype=np.float32)
    lib().orc_three_interpolate_stack(idx.shape[0], features.shape[1], _fp(features), _ip(idx), _fp(weight), _fp(out))
    return out


def stack_three_interpolate_grad(grad_out, idx, weight, M):
    """ThreeInterpolate.backward (:286-300): grad_out (N,C) * weight scattered into (M,C)."""
    grad_out, idx, weight = _f32(grad_out), _i32(idx), _f32(weight)
    g = np.zeros((M, grad_out.shape[1]), dtype=np.float32)
    for j in range(3):
        np.add.at(g, idx[:, j], grad_out * weight[:, j:j + 1])
    return g


def load_ref(name):
    """Import oracle/_ref/<name>.so (a torch pybind module built from /root/reference sources)."""
    import torch  # noqa: F401  (the extension links libtorch)
    path = os.path.join(HERE, "_ref", name + ".so")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod
