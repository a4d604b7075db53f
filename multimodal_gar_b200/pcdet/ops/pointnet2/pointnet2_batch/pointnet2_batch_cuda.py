"""Stand-in for the reference's `pointnet2_batch_cuda` pybind module
(/root/reference/pcdet/ops/pointnet2/pointnet2_batch/src/pointnet2_api.cpp:11-23): the nine wrappers with the
reference's names, argument order and caller-allocated outputs, on libmgar_lidar.so."""
import torch

from ..... import _lib
from ....._torch_glue import need_cuda, need_dtype, on_device, ptr, stream_of


def _call(t, fn, what, *args):
    with on_device(t):
        _lib.check(getattr(_lib.lib(), fn)(*args, stream_of(t)), what)
    return 1


def ball_query_wrapper(b, n, m, radius, nsample, new_xyz, xyz, idx):
    """ball_query.cpp:22-38.  new_xyz (B,M,3), xyz (B,N,3) f32, idx (B,M,nsample) i32 zero-filled: rows without a
    neighbour are left untouched (no -1 marker in this module)."""
    need_cuda(new_xyz=new_xyz, xyz=xyz, idx=idx)
    need_dtype(torch.float32, new_xyz=new_xyz, xyz=xyz)
    need_dtype(torch.int32, idx=idx)
    if b == 0 or m == 0:
        return 1
    return _call(xyz, "mgar_ball_query_stack", "ball_query_wrapper", b, b * m, float(radius), nsample, ptr(new_xyz), 0,
                 ptr(xyz), 3, 0, n, m, ptr(idx), 2, 0, 0) if n > 0 else 1


def group_points_wrapper(b, c, n, npoints, nsample, points, idx, out):
    """group_points.cpp:39-54.  points (B,C,N), idx (B,npoints,nsample) -> out (B,C,npoints,nsample)."""
    need_cuda(points=points, idx=idx, out=out)
    need_dtype(torch.float32, points=points, out=out)
    need_dtype(torch.int32, idx=idx)
    return _call(points, "mgar_batch_group_points", "group_points_wrapper", b, c, n, npoints, nsample, ptr(points), ptr(idx),
                 ptr(out))


def group_points_grad_wrapper(b, c, n, npoints, nsample, grad_out, idx, grad_points):
    """group_points.cpp:21-36.  grad_out (B,C,npoints,nsample) -> accumulated into zero-filled grad_points (B,C,N)."""
    need_cuda(grad_out=grad_out, idx=idx, grad_points=grad_points)
    need_dtype(torch.float32, grad_out=grad_out, grad_points=grad_points)
    need_dtype(torch.int32, idx=idx)
    return _call(grad_out, "mgar_batch_group_points_grad", "group_points_grad_wrapper", b, c, n, npoints, nsample,
                 ptr(grad_out), ptr(idx), ptr(grad_points))


def gather_points_wrapper(b, c, n, npoints, points, idx, out):
    """sampling.cpp:18-30.  points (B,C,N), idx (B,npoints) -> out (B,C,npoints)."""
    return group_points_wrapper(b, c, n, npoints, 1, points, idx, out)


def gather_points_grad_wrapper(b, c, n, npoints, grad_out, idx, grad_points):
    """sampling.cpp:33-46.  grad_out (B,C,npoints) -> accumulated into zero-filled grad_points (B,C,N)."""
    return group_points_grad_wrapper(b, c, n, npoints, 1, grad_out, idx, grad_points)


def farthest_point_sampling_wrapper(b, n, m, points, temp, idx):
    """sampling.cpp:49-59.  points (B,N,3), temp (B,N) filled with 1e10 (clobbered) -> idx (B,m) i32."""
    need_cuda(points=points, temp=temp, idx=idx)
    need_dtype(torch.float32, points=points, temp=temp)
    need_dtype(torch.int32, idx=idx)
    return _call(points, "mgar_fps_batch", "farthest_point_sampling_wrapper", b, n, m, ptr(points), 3, ptr(temp), ptr(idx), 0)


def three_nn_wrapper(b, n, m, unknown, known, dist2, idx):
    """interpolate.cpp:20-34.  unknown (B,N,3), known (B,M,3) -> dist2 (B,N,3) squared, idx (B,N,3)."""
    need_cuda(unknown=unknown, known=known, dist2=dist2, idx=idx)
    need_dtype(torch.float32, unknown=unknown, known=known, dist2=dist2)
    need_dtype(torch.int32, idx=idx)
    return _call(unknown, "mgar_three_nn", "three_nn_wrapper", b, n, m, ptr(unknown), ptr(known), ptr(dist2), ptr(idx))


def three_interpolate_wrapper(b, c, m, n, points, idx, weight, out):
    """interpolate.cpp:37-54.  points (B,C,M), idx / weight (B,N,3) -> out (B,C,N)."""
    need_cuda(points=points, idx=idx, weight=weight, out=out)
    need_dtype(torch.float32, points=points, weight=weight, out=out)
    need_dtype(torch.int32, idx=idx)
    return _call(points, "mgar_three_interpolate", "three_interpolate_wrapper", b, c, m, n, ptr(points), ptr(idx),
                 ptr(weight), ptr(out))


def three_interpolate_grad_wrapper(b, c, n, m, grad_out, idx, weight, grad_points):
    """interpolate.cpp:57-73.  grad_out (B,C,N) -> accumulated into zero-filled grad_points (B,C,M)."""
    need_cuda(grad_out=grad_out, idx=idx, weight=weight, grad_points=grad_points)
    need_dtype(torch.float32, grad_out=grad_out, weight=weight, grad_points=grad_points)
    need_dtype(torch.int32, idx=idx)
    return _call(grad_out, "mgar_three_interpolate_grad", "three_interpolate_grad_wrapper", b, c, n, m, ptr(grad_out),
                 ptr(idx), ptr(weight), ptr(grad_points))
