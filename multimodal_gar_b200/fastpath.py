"""Thin wrappers over the C-ABI extensions that have no counterpart in the reference's pybind surface:
strided inputs (reading the xyz columns of a pooled (.., 3+C) array in place), the uniform stack layout
(every person S points / P centroids: no count tensors, no prefix kernels) and fused outputs (FPS also
gathers the sampled coordinates).  Used by lidar_branch.PointPersonEncoder; the drop-in pcdet.ops
mirrors do not need them."""
import torch

from . import _lib
from ._torch_glue import need_cuda, need_dtype, on_device, ptr, stream_of


def _rows(t):
    """(rows, cols>=3) fp32 CUDA view with unit column stride -> row stride in floats."""
    if not t.is_cuda or t.dtype != torch.float32 or t.stride(-1) != 1:
        raise _lib.MgarError("expected a CUDA fp32 tensor with unit column stride")
    return t.stride(-2)


def fps_gather(xyz, npoint):
    """xyz (b,n,3) - possibly a column view of a (b,n,3+C) array - -> idx (b,npoint) int32, new_xyz (b,npoint,3).
    Same indices as pointnet2_utils.farthest_point_sample (reference tie order); no temp tensor for n <= 8192."""
    b, n = xyz.shape[0], xyz.shape[1]
    xs = _rows(xyz)
    if b > 1 and xyz.stride(0) != n * xs:
        raise _lib.MgarError("clouds must be densely stacked")
    idx = torch.empty((b, npoint), dtype=torch.int32, device=xyz.device)
    new_xyz = torch.empty((b, npoint, 3), dtype=torch.float32, device=xyz.device)
    temp = torch.full((b, n), 1e10, dtype=torch.float32, device=xyz.device) if n > 8192 else None
    with on_device(xyz):
        _lib.check(_lib.lib().mgar_fps_batch(b, n, npoint, ptr(xyz), xs, ptr(temp), ptr(idx), ptr(new_xyz),
                                             stream_of(xyz)), "fps_gather")
    return idx, new_xyz


def fps_ball_query(xyz, npoint, radius, nsample, all_rows=True, cloud_weight=None):
    """FPS + ball query around the sampled points in one kernel, for many equally sized power-of-two clouds.
    xyz (b,n,3) (column view of a (b,n,3+C) array allowed) -> (idx (b,npoint) int32, new_xyz (b,npoint,3),
    ball_idx (b*npoint,nsample) int32, live_cnt (b) int32) - the results of fps_gather + ball_query_uniform, plus the
    number of distinct samples per cloud (the centroids behind it repeat centroid 0) - or None when the shape is
    not covered (the caller then runs the two kernels).  all_rows=False: the ball rows of the repeated centroids are
    not written (only up to the next multiple of four) - for the consumer that walks the distinct centroids.
    cloud_weight (b) int32: distinct points per cloud (roipoint_pool_from_masks' count) - schedules the long clouds
    first; results do not depend on it."""
    b, n = xyz.shape[0], xyz.shape[1]
    L = _lib.lib()
    if not L.mgar_fps_ball_query_supported(b, n, int(npoint), int(nsample)):
        return None
    xs = _rows(xyz)
    if b > 1 and xyz.stride(0) != n * xs:
        raise _lib.MgarError("clouds must be densely stacked")
    idx = torch.empty((b, npoint), dtype=torch.int32, device=xyz.device)
    new_xyz = torch.empty((b, npoint, 3), dtype=torch.float32, device=xyz.device)
    ball_idx = torch.empty((b * npoint, nsample), dtype=torch.int32, device=xyz.device)
    live_cnt = torch.empty((b,), dtype=torch.int32, device=xyz.device)
    ws = ws_bytes = None
    if cloud_weight is not None:
        need_cuda(cloud_weight=cloud_weight)
        need_dtype(torch.int32, cloud_weight=cloud_weight)
        ws_bytes = int(L.mgar_fps_ball_query_workspace_bytes(b))
        ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=xyz.device)
    with on_device(xyz):
        _lib.check(L.mgar_fps_ball_query(b, n, int(npoint), ptr(xyz), xs, float(radius), int(nsample), ptr(idx), ptr(new_xyz),
                                         ptr(ball_idx), ptr(live_cnt), 1 if all_rows else 0, ptr(cloud_weight), ptr(ws),
                                         ws_bytes or 0, stream_of(xyz)), "fps_ball_query")
    return idx, new_xyz, ball_idx, live_cnt


def ball_query_uniform(radius, nsample, xyz, new_xyz, n, m):
    """xyz (B*n, 3) (row-strided view allowed), new_xyz (B*m, 3) contiguous -> raw idx (B*m, nsample) int32 with
    the -1 sentinel in slot 0 of empty balls (every slot written)."""
    need_cuda(new_xyz=new_xyz)
    need_dtype(torch.float32, new_xyz=new_xyz)
    xs = _rows(xyz)
    M = new_xyz.shape[0]
    B = M // m
    if B * m != M or xyz.shape[0] != B * n:
        raise _lib.MgarError("uniform layout: expected B*n points and B*m centroids")
    idx = torch.empty((M, nsample), dtype=torch.int32, device=xyz.device)
    with on_device(xyz):
        _lib.check(_lib.lib().mgar_ball_query_stack(B, M, float(radius), int(nsample), ptr(new_xyz), 0, ptr(xyz), xs, 0,
                                                    int(n), int(m), ptr(idx), 1, 0, 0, stream_of(xyz)), "ball_query_uniform")
    return idx


def box_masks(points, boxes, pool_boxes=None, want_box_idx=True):
    """points (B,N,3), boxes (B,M,7) -> (box_idx (B,N) int32 first-hit or None, mask (B,M,ceil(N/32)) int32 bit mask of
    the points inside each pooling box).  One pass over the points for all boxes; pool_boxes = enlarged boxes when
    the pooling boxes differ from the detection boxes."""
    need_cuda(points=points, boxes=boxes, pool_boxes=pool_boxes)
    need_dtype(torch.float32, points=points, boxes=boxes, pool_boxes=pool_boxes)
    B, N, M = points.shape[0], points.shape[1], boxes.shape[1]
    pool_boxes = boxes if pool_boxes is None else pool_boxes
    box_idx = torch.empty((B, N), dtype=torch.int32, device=points.device) if want_box_idx else None
    mask = torch.empty((B, M, (N + 31) // 32), dtype=torch.int32, device=points.device)
    nb = int(_lib.lib().mgar_box_masks_workspace_bytes(B, M))
    ws = torch.empty((nb,), dtype=torch.uint8, device=points.device) if nb else None     # per-frame box grids
    with on_device(points):
        _lib.check(_lib.lib().mgar_box_masks(ptr(boxes), ptr(pool_boxes), ptr(points), ptr(box_idx), ptr(mask), B, M, N,
                                             ptr(ws), nb, stream_of(points)), "box_masks")
    return box_idx, mask


def roipoint_pool_from_masks(points, point_features, pool_boxes, mask, num_sampled_points, canonical=True, want_cnt=False):
    """-> pooled (B,M,S,3+C) (canonical frame of each box when `canonical`), empty (B,M) int32; same result as
    roipoint_pool3d_cuda.forward on the same (enlarged) boxes.  want_cnt: also (B,M) int32 min(in-box points, S)."""
    need_cuda(points=points, point_features=point_features, pool_boxes=pool_boxes, mask=mask)
    B, N, M, C, S = points.shape[0], points.shape[1], pool_boxes.shape[1], point_features.shape[2], int(num_sampled_points)
    pooled = torch.empty((B, M, S, 3 + C), dtype=torch.float32, device=points.device)
    empty = torch.empty((B, M), dtype=torch.int32, device=points.device)
    cnt = torch.empty((B, M), dtype=torch.int32, device=points.device) if want_cnt else None
    with on_device(points):
        _lib.check(_lib.lib().mgar_roipoint_pool_from_masks(ptr(points), ptr(pool_boxes), ptr(point_features), ptr(mask),
                                                            ptr(pooled), ptr(empty), 0, ptr(cnt), B, N, M, S, C,
                                                            1 if canonical else 0, stream_of(points)), "roipoint_pool_from_masks")
    return (pooled, empty, cnt) if want_cnt else (pooled, empty)
