"""The C-ABI library loads and exports every symbol include/mgar_lidar.h declares (no compute calls)."""
import ctypes
import os
import re

import pytest

from multimodal_gar_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "mgar_lidar.h")


def declared_symbols():
    text = open(HEADER).read()
    return sorted(set(re.findall(r"MGAR_API\s+[\w\s\*]+?\b(mgar_\w+)\s*\(", text)))


def test_header_declares_the_hot_path():
    names = declared_symbols()
    for need in ("mgar_points_in_boxes", "mgar_points_in_boxes_mask_host", "mgar_roipoint_pool", "mgar_fps_batch",
                 "mgar_fps_stack", "mgar_ball_query_stack", "mgar_group_points_stack",
                 "mgar_group_points_grad_stack", "mgar_sa_mlp_fwd_fp32"):
        assert need in names


def test_library_exports_every_declared_symbol():
    assert os.path.exists(_lib.LIB_PATH), "libmgar_lidar.so missing: run __graft_entry__.build()"
    L = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(L, name), "symbol %s declared in the header but not exported" % name


def test_ctypes_table_matches_header():
    assert sorted(_lib.SIGNATURES) == declared_symbols()
    text = open(HEADER).read()
    for name, args in _lib.SIGNATURES.items():
        m = re.search(r"MGAR_API[^;(]*\b%s\s*\(([^;]*?)\)\s*;" % name, text, re.S)
        assert m, name
        params = m.group(1).strip()
        n = 0 if params in ("void", "") else len(params.split(","))
        assert n == len(args), "%s: header has %d parameters, ctypes table %d" % (name, n, len(args))


def test_abi_version_and_error_string():
    L = _lib.lib()
    assert L.mgar_abi_version() == 2
    assert isinstance(L.mgar_last_error(), bytes)


def test_bad_arguments_fail_without_a_gpu():
    # argument validation happens before any CUDA call: negative sizes are rejected with a message
    L = _lib.lib()
    rc = L.mgar_points_in_boxes(0, 0, 0, -1, 1, 1, 0)
    assert rc == -1 and b"negative" in L.mgar_last_error()
    with pytest.raises(_lib.MgarError):
        _lib.check(rc, "points_in_boxes")


def test_ops_refuse_cpu_tensors():
    """No CPU fallback: the GPU entry points raise on host tensors instead of computing."""
    import torch
    from multimodal_gar_b200.pcdet.ops.roiaware_pool3d import roiaware_pool3d_utils
    with pytest.raises(_lib.MgarError):
        roiaware_pool3d_utils.points_in_boxes_gpu(torch.zeros(1, 8, 3), torch.zeros(1, 2, 7))
