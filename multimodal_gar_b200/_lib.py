"""ctypes face of libmgar_lidar.so (the C ABI declared in include/mgar_lidar.h).

The library is the product: if it is missing or fails to load, importing any op raises - there
is no CPU or PyTorch fallback behind these calls.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MGAR_LIB", os.path.join(_HERE, "libmgar_lidar.so"))   # MGAR_LIB: experiment builds only
_lib = None

c_fp = ctypes.c_void_p  # device/host pointers travel as integers (tensor.data_ptr())
_i = ctypes.c_int
_f = ctypes.c_float
_p = ctypes.c_void_p

# name -> argtypes; mirrors include/mgar_lidar.h one to one (tests check the export list).
SIGNATURES = {
    "mgar_abi_version": [],
    "mgar_last_error": [],
    "mgar_launch_count": [],
    "mgar_points_in_boxes": [_p, _p, _p, _i, _i, _i, _p],
    "mgar_in_box_mask": [_p, _p, _p, _p, _i, _i, _i, _p],
    "mgar_points_in_boxes_mask_host_workspace_bytes": [_i, _i],
    "mgar_points_in_boxes_mask_host": [_p, _p, _p, _i, _i, _p, ctypes.c_size_t],
    "mgar_roipoint_pool": [_p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p],
    "mgar_box_mask_words": [_i],
    "mgar_box_masks_workspace_bytes": [_i, _i],
    "mgar_box_masks": [_p, _p, _p, _p, _p, _i, _i, _i, _p, ctypes.c_size_t, _p],
    "mgar_roipoint_pool_from_masks": [_p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p],
    "mgar_roiaware_pool3d_workspace_bytes": [_i, _i],
    "mgar_roiaware_pool3d_fwd": [_p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _i, _i, _p],
    "mgar_roiaware_pool3d_bwd": [_p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _i, _p],
    "mgar_fps_batch": [_i, _i, _i, _p, _i, _p, _p, _p, _p],
    "mgar_fps_stack": [_i, _i, _p, _p, _p, _p, _p, _p],
    "mgar_stack_workspace_bytes": [_i],
    "mgar_host_alloc": [_p, ctypes.c_size_t, _i],
    "mgar_host_free": [_p],
    "mgar_ball_query_stack": [_i, _i, _f, _i, _p, _p, _p, _i, _p, _i, _i, _p, _i, _p, ctypes.c_size_t, _p],
    "mgar_voxel_query_stack": [_i, _i, _i, _i, _i, _f, _i, _i, _i, _p, _p, _p, _p, _p, _p],
    "mgar_batch_group_points": [_i, _i, _i, _i, _i, _p, _p, _p, _p],
    "mgar_batch_group_points_grad": [_i, _i, _i, _i, _i, _p, _p, _p, _p],
    "mgar_three_nn": [_i, _i, _i, _p, _p, _p, _p, _p],
    "mgar_three_interpolate": [_i, _i, _i, _i, _p, _p, _p, _p, _p],
    "mgar_three_interpolate_grad": [_i, _i, _i, _i, _p, _p, _p, _p, _p],
    "mgar_three_nn_stack": [_i, _i, _i, _p, _p, _p, _p, _p, _p, _p, ctypes.c_size_t, _p],
    "mgar_three_interpolate_stack": [_i, _i, _p, _p, _p, _p, _p],
    "mgar_three_interpolate_grad_stack": [_i, _i, _p, _p, _p, _p, _p],
    "mgar_dafm_attention": [_i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _f, _p, _p],
    "mgar_dafm_add_layernorm": [_i, _i, _p, _p, _p, _p, _p, _f, _p, _p],
    "mgar_voxel_sa_supported": [_i, _i, _i, _i],
    "mgar_voxel_sa_param_floats": [_i, _i, _i],
    "mgar_voxel_sa_workspace_bytes": [_i, _i],
    "mgar_voxel_sa_fwd": [_i, _i, _i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _i, _p, ctypes.c_size_t, _p],
    "mgar_voxel_sa_fwd_static": [_i, _i, _i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _i, _p],
    "mgar_gemm_tf32x3_supported": [_i, _i, _i],
    "mgar_gemm_tf32x3": [_i, _i, _i, _p, _i, _p, _i, _p, _i, _p, _i, _p],
    "mgar_gemm_packed_bytes": [_i, _i],
    "mgar_gemm_pack_bt": [_i, _i, _p, _i, _p, _p],
    "mgar_gemm_tf32x3_packed": [_i, _i, _i, _p, _i, _p, _p, _i, _p, _i, _p],
    "mgar_gemm_tf32x3_packed_grouped": [_i, _i, _i, _i, _p, _p, _p, _p, _i, _p, _p, _p],
    "mgar_fps_ball_query_supported": [_i, _i, _i, _i],
    "mgar_fps_ball_query_workspace_bytes": [_i],
    "mgar_fps_ball_query": [_i, _i, _i, _p, _i, _f, _i, _p, _p, _p, _p, _i, _p, _p, ctypes.c_size_t, _p],
    "mgar_group_points_stack": [_i, _i, _i, _i, _p, _p, _p, _p, _p, _p, ctypes.c_size_t, _p],
    "mgar_group_points_grad_stack": [_i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, ctypes.c_size_t, _p],
    "mgar_sa_mlp_tc_supported": [_i, _i, _i, _p, _i],
    "mgar_sa_mlp_tc_group_max_supported": [_i, _i, _i, _p, _i],
    "mgar_sa_mlp_tc_workspace_bytes": [_i, _i],
    "mgar_sa_mlp_tc_unit_slots": [],
    "mgar_sa_mlp_fwd_tc": [_i, _i, _i, _i, _p, _i, _p, _p, _p, _i, _i, _p, _i, _p, _p, _i, _p, _p, _p, _p, _i, _p, _i, _p, _p,
                           _p, _i, _p, ctypes.c_size_t, _p],
    "mgar_sa_mlp_fwd_fp32": [_i, _i, _i, _i, _p, _p, _p, _p, _i, _i, _p, _p, _p, _i, _p, _p, _p, _i, _i, _p, ctypes.c_size_t, _p],
    "mgar_sa_mlp_train_workspace_bytes": [_i, _i, _i, _p],
    "mgar_sa_mlp_train_workspace_bytes_for": [_i, _i, _i, _p, _i],
    "mgar_sa_mlp_fwd_train": [_i, _i, _i, _i, _p, _p, _p, _p, _i, _i, _p, _p, _p, _i, _p, _p, _p, _f, _p, _i, _i, _p, _p,
                              ctypes.c_size_t, _p],
    "mgar_sa_mlp_bwd": [_i, _i, _i, _i, _p, _p, _i, _i, _p, _p, _i, _p, _p, _p, _p, _i, _i, _p, ctypes.c_size_t, _p, _p, _p, _p],
}


class MgarError(RuntimeError):
    pass


def build(verbose=False):
    """Compile libmgar_lidar.so in-tree with nvcc for sm_100a (no GPU needed)."""
    cmd = ["make", "-C", os.path.join(_HERE, "csrc"), "-j8"]
    out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if out.returncode != 0 or verbose:
        print(out.stdout)
    if out.returncode != 0:
        raise MgarError("building libmgar_lidar.so failed")


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MgarError(
                "libmgar_lidar.so is not built (%s): run `python -c 'import __graft_entry__ as g; g.build()'`; "
                "there is no fallback path" % LIB_PATH)
        L = ctypes.CDLL(LIB_PATH)
        for name, args in SIGNATURES.items():
            fn = getattr(L, name)
            fn.argtypes = args
            fn.restype = _i
        L.mgar_last_error.restype = ctypes.c_char_p
        L.mgar_launch_count.restype = ctypes.c_ulonglong
        L.mgar_sa_mlp_train_workspace_bytes.restype = ctypes.c_size_t
        L.mgar_sa_mlp_train_workspace_bytes_for.restype = ctypes.c_size_t
        L.mgar_box_mask_words.restype = ctypes.c_size_t
        L.mgar_stack_workspace_bytes.restype = ctypes.c_size_t
        L.mgar_gemm_packed_bytes.restype = ctypes.c_size_t
        L.mgar_voxel_sa_param_floats.restype = ctypes.c_size_t
        L.mgar_voxel_sa_workspace_bytes.restype = ctypes.c_size_t
        L.mgar_box_masks_workspace_bytes.restype = ctypes.c_size_t
        L.mgar_points_in_boxes_mask_host_workspace_bytes.restype = ctypes.c_size_t
        L.mgar_sa_mlp_tc_workspace_bytes.restype = ctypes.c_size_t
        L.mgar_fps_ball_query_workspace_bytes.restype = ctypes.c_size_t
        L.mgar_roiaware_pool3d_workspace_bytes.restype = ctypes.c_size_t
        _lib = L
    return _lib


def check(rc, what=""):
    if rc != 0:
        msg = lib().mgar_last_error().decode("utf-8", "replace")
        raise MgarError("%s failed (code %d): %s" % (what or "libmgar_lidar call", rc, msg))


def stack_workspace(B, device):
    """Device scratch of the stack-layout operators (mgar_stack_workspace_bytes(B) bytes from torch's caching allocator:
    the library itself allocates nothing) -> (tensor, nbytes)."""
    import torch
    n = int(lib().mgar_stack_workspace_bytes(int(B)))
    return torch.empty((max(n, 16),), dtype=torch.uint8, device=device), n


def launch_count():
    return int(lib().mgar_launch_count())
