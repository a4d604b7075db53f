"""Stand-in for the reference's `pointnet2_stack_cuda` pybind module
(/root/reference/pcdet/ops/pointnet2/pointnet2_stack/src/pointnet2_api.cpp:12-31): the five
hot-path wrappers, voxel_query_wrapper and the three feature-propagation wrappers (three_nn / three_interpolate /
three_interpolate_grad) run on libmgar_lidar.so; the four vector-pool names exist (so imports succeed) and raise
NotImplementedError - they belong to PV-RCNN++ paths outside this build.
"""
import torch

from ..... import _lib
from ....._torch_glue import need_cuda, need_dtype, on_device, ptr, stream_of


def ball_query_wrapper(B, M, radius, nsample, new_xyz, new_xyz_batch_cnt, xyz, xyz_batch_cnt, idx):
    """ball_query.cpp:29-45. idx (M,nsample) i32 zero-filled by the caller."""
    need_cuda(new_xyz=new_xyz, new_xyz_batch_cnt=new_xyz_batch_cnt, xyz=xyz, xyz_batch_cnt=xyz_batch_cnt, idx=idx)
    need_dtype(torch.float32, new_xyz=new_xyz, xyz=xyz)
    need_dtype(torch.int32, new_xyz_batch_cnt=new_xyz_batch_cnt, xyz_batch_cnt=xyz_batch_cnt, idx=idx)
    with on_device(xyz):
        ws, nb = _lib.stack_workspace(B, xyz.device)      # the binding owns the scratch (torch's allocator), SURVEY 8b
        _lib.check(_lib.lib().mgar_ball_query_stack(int(B), int(M), float(radius), int(nsample), ptr(new_xyz),
                                                    ptr(new_xyz_batch_cnt), ptr(xyz), 3, ptr(xyz_batch_cnt), 0, 0,
                                                    ptr(idx), 0, ptr(ws), nb, stream_of(xyz)), "ball_query_wrapper")
    return 1


def group_points_wrapper(B, M, C, nsample, features, features_batch_cnt, idx, idx_batch_cnt, out):
    """group_points.cpp:50-67. out (M,C,nsample) f32, every element written."""
    need_cuda(features=features, features_batch_cnt=features_batch_cnt, idx=idx, idx_batch_cnt=idx_batch_cnt, out=out)
    need_dtype(torch.float32, features=features, out=out)
    need_dtype(torch.int32, features_batch_cnt=features_batch_cnt, idx=idx, idx_batch_cnt=idx_batch_cnt)
    with on_device(features):
        ws, nb = _lib.stack_workspace(B, features.device)
        _lib.check(_lib.lib().mgar_group_points_stack(int(B), int(M), int(C), int(nsample), ptr(features),
                                                      ptr(features_batch_cnt), ptr(idx), ptr(idx_batch_cnt), ptr(out),
                                                      ptr(ws), nb, stream_of(features)), "group_points_wrapper")
    return 1


def group_points_grad_wrapper(B, M, C, N, nsample, grad_out, idx, idx_batch_cnt, features_batch_cnt, grad_features):
    """group_points.cpp:29-47. grad_features (N,C) zero-filled by the caller."""
    need_cuda(grad_out=grad_out, idx=idx, idx_batch_cnt=idx_batch_cnt, features_batch_cnt=features_batch_cnt,
              grad_features=grad_features)
    need_dtype(torch.float32, grad_out=grad_out, grad_features=grad_features)
    need_dtype(torch.int32, idx=idx, idx_batch_cnt=idx_batch_cnt, features_batch_cnt=features_batch_cnt)
    with on_device(grad_out):
        ws, nb = _lib.stack_workspace(B, grad_out.device)
        _lib.check(_lib.lib().mgar_group_points_grad_stack(int(B), int(M), int(C), int(N), int(nsample), ptr(grad_out),
                                                           ptr(idx), ptr(idx_batch_cnt), ptr(features_batch_cnt),
                                                           ptr(grad_features), ptr(ws), nb, stream_of(grad_out)),
                   "group_points_grad_wrapper")
    return 1


def farthest_point_sampling_wrapper(b, n, m, points, temp, idx):
    """sampling.cpp:21-34. points (b,n,3), temp (b,n) pre-filled 1e10 (clobbered), idx (b,m) i32."""
    need_cuda(points=points, temp=temp, idx=idx)
    need_dtype(torch.float32, points=points, temp=temp)
    need_dtype(torch.int32, idx=idx)
    with on_device(points):
        _lib.check(_lib.lib().mgar_fps_batch(int(b), int(n), int(m), ptr(points), 3, ptr(temp), ptr(idx), 0,
                                             stream_of(points)), "farthest_point_sampling_wrapper")
    return 1


def stack_farthest_point_sampling_wrapper(points, temp, xyz_batch_cnt, idx, num_sampled_points):
    """sampling.cpp:37-56. points (N,3), temp (N), xyz_batch_cnt (B) i32, idx (sum m) i32, num_sampled_points (B) i32."""
    need_cuda(points=points, temp=temp, xyz_batch_cnt=xyz_batch_cnt, idx=idx, num_sampled_points=num_sampled_points)
    need_dtype(torch.float32, points=points, temp=temp)
    need_dtype(torch.int32, xyz_batch_cnt=xyz_batch_cnt, idx=idx, num_sampled_points=num_sampled_points)
    batch = xyz_batch_cnt.numel()
    # Tight per-cloud bound keeps clouds in registers; one host read (the reference's Python wrapper
    # already synchronises on npoint.sum().item(), pointnet2_utils.py:217).
    max_n = int(xyz_batch_cnt.max().item()) if batch > 0 else 0
    with on_device(points):
        _lib.check(_lib.lib().mgar_fps_stack(batch, max_n, ptr(points), ptr(temp), ptr(xyz_batch_cnt), ptr(idx),
                                             ptr(num_sampled_points), stream_of(points)),
                   "stack_farthest_point_sampling_wrapper")
    return 1


def _outside(name):
    def fn(*args, **kwargs):
        raise NotImplementedError("%s is outside the LiDAR hot path of this build (SURVEY.md 2.3 / 8f)" % name)
    fn.__name__ = name
    return fn


def voxel_query_wrapper(M, R1, R2, R3, nsample, radius, z_range, y_range, x_range, new_xyz, xyz, new_coords,
                        point_indices, idx):
    """voxel_query.cpp:25-41.  new_xyz (M,3) f32, xyz (N,3) f32, new_coords (M,4) i32 [batch,z,y,x],
    point_indices (B,R1,R2,R3) i32, idx (M,nsample) i32 zero-filled by the caller.  Returns 1."""
    need_cuda(new_xyz=new_xyz, xyz=xyz, new_coords=new_coords, point_indices=point_indices, idx=idx)
    need_dtype(torch.float32, new_xyz=new_xyz, xyz=xyz)
    need_dtype(torch.int32, new_coords=new_coords, point_indices=point_indices, idx=idx)
    with on_device(new_xyz):
        _lib.check(_lib.lib().mgar_voxel_query_stack(M, R1, R2, R3, nsample, float(radius), z_range, y_range, x_range,
                                                     ptr(new_xyz), ptr(xyz), ptr(new_coords), ptr(point_indices), ptr(idx),
                                                     stream_of(new_xyz)), "voxel_query_wrapper")
    return 1


def three_nn_wrapper(unknown, unknown_batch_cnt, known, known_batch_cnt, dist2, idx):
    """interpolate.cpp:32-60.  unknown (N,3), known (M,3) f32; *_batch_cnt (B) i32; dist2 (N,3) f32 (squared) and
    idx (N,3) i32 (global known rows) are filled in place."""
    need_cuda(unknown=unknown, unknown_batch_cnt=unknown_batch_cnt, known=known, known_batch_cnt=known_batch_cnt,
              dist2=dist2, idx=idx)
    need_dtype(torch.float32, unknown=unknown, known=known, dist2=dist2)
    need_dtype(torch.int32, unknown_batch_cnt=unknown_batch_cnt, known_batch_cnt=known_batch_cnt, idx=idx)
    with on_device(unknown):
        ws, nb = _lib.stack_workspace(unknown_batch_cnt.numel(), unknown.device)
        _lib.check(_lib.lib().mgar_three_nn_stack(unknown_batch_cnt.numel(), unknown.shape[0], known.shape[0], ptr(unknown),
                                                  ptr(unknown_batch_cnt), ptr(known), ptr(known_batch_cnt), ptr(dist2),
                                                  ptr(idx), ptr(ws), nb, stream_of(unknown)), "three_nn_wrapper")


def three_interpolate_wrapper(features, idx, weight, out):
    """interpolate.cpp:63-83.  features (M,C) f32, idx (N,3) i32, weight (N,3) f32 -> out (N,C) f32."""
    need_cuda(features=features, idx=idx, weight=weight, out=out)
    need_dtype(torch.float32, features=features, weight=weight, out=out)
    need_dtype(torch.int32, idx=idx)
    with on_device(features):
        _lib.check(_lib.lib().mgar_three_interpolate_stack(idx.shape[0], features.shape[1], ptr(features), ptr(idx),
                                                           ptr(weight), ptr(out), stream_of(features)),
                   "three_interpolate_wrapper")


def three_interpolate_grad_wrapper(grad_out, idx, weight, grad_features):
    """interpolate.cpp:86-107.  grad_out (N,C) f32 scattered into zero-filled grad_features (M,C)."""
    need_cuda(grad_out=grad_out, idx=idx, weight=weight, grad_features=grad_features)
    need_dtype(torch.float32, grad_out=grad_out, weight=weight, grad_features=grad_features)
    need_dtype(torch.int32, idx=idx)
    with on_device(grad_out):
        _lib.check(_lib.lib().mgar_three_interpolate_grad_stack(grad_out.shape[0], grad_out.shape[1], ptr(grad_out),
                                                                ptr(idx), ptr(weight), ptr(grad_features),
                                                                stream_of(grad_out)), "three_interpolate_grad_wrapper")


query_stacked_local_neighbor_idxs_wrapper_stack = _outside("query_stacked_local_neighbor_idxs_wrapper_stack")
query_three_nn_by_stacked_local_idxs_wrapper_stack = _outside("query_three_nn_by_stacked_local_idxs_wrapper_stack")
vector_pool_wrapper = _outside("vector_pool_wrapper")
vector_pool_grad_wrapper = _outside("vector_pool_grad_wrapper")
