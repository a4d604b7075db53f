"""The two helpers of the reference's pcdet/utils/common_utils.py that the LiDAR hot path uses."""
import numpy as np
import torch


def check_numpy_to_torch(x):
    """pcdet/utils/common_utils.py:19-24: numpy arrays become float tensors, with a flag to convert back."""
    if isinstance(x, np.ndarray):
        return torch.from_numpy(x).float(), True
    if isinstance(x, (int, float)):
        return torch.tensor([x]).float(), True
    return x, False


def rotate_points_along_z(points, angle):
    """pcdet/utils/common_utils.py:35-57: points (B,N,3+C), angle (B) -> rotated copy (x->y positive)."""
    points, is_numpy = check_numpy_to_torch(points)
    angle, _ = check_numpy_to_torch(angle)
    c, s = torch.cos(angle), torch.sin(angle)
    o, z = torch.ones_like(angle), torch.zeros_like(angle)
    rot = torch.stack((c, s, z, -s, c, z, z, z, o), dim=1).view(-1, 3, 3).float()
    out = torch.cat((torch.matmul(points[:, :, 0:3], rot), points[:, :, 3:]), dim=-1)
    return out.numpy() if is_numpy else out


def scatter_point_inds(indices, point_inds, shape):
    """pcdet/utils/common_utils.py:235-241: dense tensor of `shape` holding point_inds at `indices`, -1 elsewhere."""
    ret = torch.full(tuple(shape), -1, dtype=point_inds.dtype, device=point_inds.device)
    flat = indices.view(-1, indices.shape[-1])
    ret[tuple(flat[:, i] for i in range(flat.shape[1]))] = point_inds
    return ret


def generate_voxel2pinds(sparse_tensor):
    """pcdet/utils/common_utils.py:244-252: (B,Z,Y,X) int32 map voxel -> row of `sparse_tensor.indices` (-1 = empty).
    Accepts anything with `.indices` (V,4) [batch,z,y,x], `.batch_size`, `.spatial_shape` (a spconv tensor has them)."""
    indices = sparse_tensor.indices.long()
    rows = torch.arange(indices.shape[0], device=indices.device, dtype=torch.int32)
    return scatter_point_inds(indices, rows, [sparse_tensor.batch_size] + list(sparse_tensor.spatial_shape))


def get_voxel_centers(voxel_coords, downsample_times, voxel_size, point_cloud_range):
    """pcdet/utils/common_utils.py:66-82: voxel_coords (N,3) [z,y,x] -> centres (N,3) [x,y,z] in metres."""
    assert voxel_coords.shape[1] == 3
    centers = voxel_coords[:, [2, 1, 0]].float()
    size = torch.tensor(voxel_size, device=centers.device).float() * downsample_times
    origin = torch.tensor(point_cloud_range[0:3], device=centers.device).float()
    return (centers + 0.5) * size + origin
