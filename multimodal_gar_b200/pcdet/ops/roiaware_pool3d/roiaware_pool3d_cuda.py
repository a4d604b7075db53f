"""Stand-in for the reference's `roiaware_pool3d_cuda` pybind module
(/root/reference/pcdet/ops/roiaware_pool3d/src/roiaware_pool3d.cpp:172-177): same function names,
argument order (boxes first, points second) and caller-allocated outputs, backed by
libmgar_lidar.so.  Dropping this file next to the reference's roiaware_pool3d_utils.py makes
that wrapper run on the new kernels unchanged.
"""
import torch

from .... import _lib
from ...._torch_glue import need_cuda, need_dtype, on_device, ptr, stream_of


def points_in_boxes_gpu(boxes, pts, box_idx_of_points):
    """boxes (B,M,7) f32 cuda, pts (B,N,3) f32 cuda, box_idx_of_points (B,N) i32 cuda. Returns 1."""
    need_cuda(boxes=boxes, pts=pts, box_idx_of_points=box_idx_of_points)
    need_dtype(torch.float32, boxes=boxes, pts=pts)
    need_dtype(torch.int32, box_idx_of_points=box_idx_of_points)
    B, M, N = boxes.size(0), boxes.size(1), pts.size(1)
    with on_device(pts):
        _lib.check(_lib.lib().mgar_points_in_boxes(ptr(boxes), ptr(pts), ptr(box_idx_of_points), B, M, N,
                                                   stream_of(pts)), "points_in_boxes_gpu")
    return 1


def points_in_boxes_cpu(boxes, pts, pts_indices):
    """boxes (M,7) f32 cpu, pts (N,3) f32 cpu, pts_indices (M,N) i32 cpu (host buffers in and out;
    the work happens on the GPU - roiaware_pool3d.cpp:143-168 semantics, margin 1e-2)."""
    for name, t in (("boxes", boxes), ("pts", pts), ("pts_indices", pts_indices)):
        if t.is_cuda or not t.is_contiguous():
            raise _lib.MgarError("%s must be a contiguous CPU tensor" % name)
    need_dtype(torch.float32, boxes=boxes, pts=pts)
    need_dtype(torch.int32, pts_indices=pts_indices)
    M, N = boxes.size(0), pts.size(0)
    L = _lib.lib()
    if not torch.cuda.is_available():
        raise _lib.MgarError("points_in_boxes_cpu computes on the GPU (host buffers in and out): no CUDA device, no fallback")
    nb = int(L.mgar_points_in_boxes_mask_host_workspace_bytes(M, N))
    # device staging on the CURRENT device, from torch's caching allocator (the library allocates nothing)
    ws = torch.empty((max(nb, 16),), dtype=torch.uint8, device=torch.device("cuda", torch.cuda.current_device()))
    _lib.check(L.mgar_points_in_boxes_mask_host(ptr(boxes), ptr(pts), ptr(pts_indices), M, N, ptr(ws), nb),
               "points_in_boxes_cpu")
    return 1


def forward(rois, pts, pts_feature, argmax, pts_idx_of_voxels, pooled_features, pool_method):
    """roiaware_pool3d_gpu (roiaware_pool3d.cpp:20-58).  rois (N,7), pts (npoints,3), pts_feature (npoints,C) f32
    cuda; argmax (N,ox,oy,oz,C) i32, pts_idx_of_voxels (N,ox,oy,oz,max_pts) i32 ZERO-FILLED, pooled_features
    (N,ox,oy,oz,C) f32 zero-filled, all caller-allocated; pool_method 0 = max, 1 = avg.  Returns 1."""
    need_cuda(rois=rois, pts=pts, pts_feature=pts_feature, argmax=argmax, pts_idx_of_voxels=pts_idx_of_voxels,
              pooled_features=pooled_features)
    need_dtype(torch.float32, rois=rois, pts=pts, pts_feature=pts_feature, pooled_features=pooled_features)
    need_dtype(torch.int32, argmax=argmax, pts_idx_of_voxels=pts_idx_of_voxels)
    boxes_num, pts_num, channels = rois.size(0), pts.size(0), pts_feature.size(1)
    max_pts = pts_idx_of_voxels.size(4)
    out_x, out_y, out_z = pts_idx_of_voxels.size(1), pts_idx_of_voxels.size(2), pts_idx_of_voxels.size(3)
    if pts_feature.size(0) != pts_num:                                       # roiaware_pool3d.cpp:45 assert
        raise _lib.MgarError("pts_feature has %d rows for %d points" % (pts_feature.size(0), pts_num))
    L = _lib.lib()
    with on_device(pts):
        nbytes = L.mgar_roiaware_pool3d_workspace_bytes(boxes_num, pts_num)
        ws = torch.empty(max(nbytes, 4) // 4, dtype=torch.int32, device=pts.device)
        _lib.check(L.mgar_roiaware_pool3d_fwd(ptr(rois), ptr(pts), ptr(pts_feature), ptr(argmax), ptr(pts_idx_of_voxels),
                                              ptr(pooled_features), ptr(ws), boxes_num, pts_num, channels, max_pts,
                                              out_x, out_y, out_z, int(pool_method), stream_of(pts)), "roiaware_pool3d forward")
    return 1


def backward(pts_idx_of_voxels, argmax, grad_out, grad_in, pool_method):
    """roiaware_pool3d_gpu_backward (roiaware_pool3d.cpp:61-96).  grad_out (N,ox,oy,oz,C) f32 -> accumulated into
    the caller's zero-filled grad_in (npoints,C).  Returns 1."""
    need_cuda(pts_idx_of_voxels=pts_idx_of_voxels, argmax=argmax, grad_out=grad_out, grad_in=grad_in)
    need_dtype(torch.float32, grad_out=grad_out, grad_in=grad_in)
    need_dtype(torch.int32, argmax=argmax, pts_idx_of_voxels=pts_idx_of_voxels)
    boxes_num, out_x, out_y, out_z, max_pts = pts_idx_of_voxels.shape
    channels = grad_out.size(4)
    with on_device(grad_out):
        _lib.check(_lib.lib().mgar_roiaware_pool3d_bwd(ptr(pts_idx_of_voxels), ptr(argmax), ptr(grad_out), ptr(grad_in),
                                                       boxes_num, out_x, out_y, out_z, channels, max_pts, int(pool_method),
                                                       stream_of(grad_out)), "roiaware_pool3d backward")
    return 1
