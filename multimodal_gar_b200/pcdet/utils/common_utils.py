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
