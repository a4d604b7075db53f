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
    _lib.check(_lib.lib().mgar_points_in_boxes_mask_host(ptr(boxes), ptr(pts), ptr(pts_indices), M, N),
               "points_in_boxes_cpu")
    return 1


def forward(*args, **kwargs):
    raise NotImplementedError("RoIAwarePool3d forward is outside the LiDAR hot path (SURVEY.md 8f-2)")


def backward(*args, **kwargs):
    raise NotImplementedError("RoIAwarePool3d backward is outside the LiDAR hot path (SURVEY.md 8f-2)")
