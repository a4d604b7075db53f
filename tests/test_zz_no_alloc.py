"""-m gpu: mgar_voxel_sa_fwd_static - the variant of mgar_voxel_sa_fwd whose scratch is a file-scope device array
(include/mgar_lidar.h; no counterpart in the reference) - against the
existing entry point with a caller-owned workspace, through the C-ABI, on two calls in a row (the second with other
sizes, so that it meets the first call's leftovers in the shared array), plus the size limit.

Collected last on purpose (test_zz_*).  Stands alone: the test builds its own inputs and
passes in a fresh process as well as after the other tests."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from multimodal_gar_b200 import _lib

pytestmark = pytest.mark.gpu
HEADER = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "mgar_lidar.h")


def _max_floats():
    m = re.search(r"#define\s+MGAR_VOXEL_SA_STATIC_MAX_FLOATS\s+\((\d+)u\s*<<\s*(\d+)\)", open(HEADER).read())
    assert m, "MGAR_VOXEL_SA_STATIC_MAX_FLOATS not found in include/mgar_lidar.h"
    return int(m.group(1)) << int(m.group(2))


def _case(seed, N, M, C_in, C1, C2, ns, dev):
    g = torch.Generator().manual_seed(seed)
    L = _lib.lib()
    n_par = int(L.mgar_voxel_sa_param_floats(C_in, C1, C2))
    t = dict(features=torch.randn((N, C_in), generator=g), xyz=torch.rand((N, 3), generator=g) * 10,
             new_xyz=torch.rand((M, 3), generator=g) * 10,
             idx=torch.randint(0, N, (M, ns), generator=g, dtype=torch.int32),
             empty=(torch.rand((M,), generator=g) < 0.1).to(torch.uint8),
             params=torch.randn((n_par,), generator=g) * 0.3)
    t["idx"][t["empty"].bool()] = 0                          # rows of empty queries are zeroed, as VoxelQuery leaves them
    return {k: v.to(dev).contiguous() for k, v in t.items()}


def _run(static, t, N, M, C_in, C1, C2, ns, pool_avg, dev):
    L = _lib.lib()
    out = torch.full((M, C2), float("nan"), device=dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    head = (N, M, C_in, C1, C2, ns, pool_avg, t["features"].data_ptr(), t["xyz"].data_ptr(), t["new_xyz"].data_ptr(),
            t["idx"].data_ptr(), t["empty"].data_ptr(), t["params"].data_ptr(), out.data_ptr(), out.stride(0))
    if static:
        rc = L.mgar_voxel_sa_fwd_static(*head, stream)
    else:
        nb = int(L.mgar_voxel_sa_workspace_bytes(N, C1))
        ws = torch.empty((max(nb, 1),), dtype=torch.uint8, device=dev)
        rc = L.mgar_voxel_sa_fwd(*head, ws.data_ptr(), nb, stream)
    _lib.check(rc, "voxel_sa_fwd_static" if static else "voxel_sa_fwd")
    torch.cuda.synchronize(dev)
    return out


def test_static_scratch_matches_workspace_entry_twice_in_a_row(cuda):
    # call 1 fills (N1, C1) rows of the shared array; call 2 is smaller and of another width: it must not see them
    shapes = [dict(seed=11, N=20000, M=7001, C_in=32, C1=32, C2=32, ns=16, pool_avg=0),
              dict(seed=12, N=3001, M=9000, C_in=16, C1=64, C2=48, ns=8, pool_avg=1)]
    with torch.cuda.device(cuda):
        for s in shapes:
            seed, pool_avg = s["seed"], s["pool_avg"]
            dims = [s[k] for k in ("N", "M", "C_in", "C1", "C2", "ns")]
            t = _case(seed, *dims, cuda)
            want = _run(False, t, *dims, pool_avg, cuda)
            got = _run(True, t, *dims, pool_avg, cuda)
            assert torch.isfinite(want).all() and torch.isfinite(got).all()
            assert torch.equal(got, want), "static-scratch entry differs from the workspace entry (%s)" % (s,)   # same kernels: same bits
        # and once more with the first shape: repeated use of the array
        s = shapes[0]
        dims = [s[k] for k in ("N", "M", "C_in", "C1", "C2", "ns")]
        t = _case(s["seed"], *dims, cuda)
        assert torch.equal(_run(True, t, *dims, s["pool_avg"], cuda), _run(False, t, *dims, s["pool_avg"], cuda))


def test_static_scratch_allocates_nothing_and_rejects_oversize(cuda):
    L = _lib.lib()
    cap = _max_floats()
    assert cap * 4 == 8 << 20
    with torch.cuda.device(cuda):
        dims = (4096, 512, 32, 32, 32, 16)
        t = _case(13, *dims, cuda)
        _run(True, t, *dims, 0, cuda)                        # set-up call (resolves the array's address)
        torch.cuda.synchronize(cuda)
        free0, _ = torch.cuda.mem_get_info(cuda)
        before = torch.cuda.memory_stats(cuda)["allocation.all.allocated"]
        out = torch.empty((dims[1], dims[4]), device=cuda)
        stream = ctypes.c_void_p(torch.cuda.current_stream(cuda).cuda_stream)
        before = torch.cuda.memory_stats(cuda)["allocation.all.allocated"]
        for _ in range(3):
            _lib.check(L.mgar_voxel_sa_fwd_static(*dims, 0, t["features"].data_ptr(), t["xyz"].data_ptr(), t["new_xyz"].data_ptr(),
                                                  t["idx"].data_ptr(), t["empty"].data_ptr(), t["params"].data_ptr(),
                                                  out.data_ptr(), out.stride(0), stream), "voxel_sa_fwd_static")
        torch.cuda.synchronize(cuda)
        assert torch.cuda.memory_stats(cuda)["allocation.all.allocated"] == before      # nothing through torch's allocator
        assert torch.cuda.mem_get_info(cuda)[0] == free0                                   # nothing through cudaMalloc either
        # one float beyond the limit: refused with a message, nothing launched
        C1 = 128
        N_big = cap // C1 + 1
        rc = L.mgar_voxel_sa_fwd_static(N_big, 1, 1, C1, 1, 1, 0, None, None, None, None, None, None, None, 1, stream)
        assert rc == -3                                                                    # MGAR_ENOMEM
        msg = L.mgar_last_error().decode()
        assert "exceeds the built-in scratch" in msg and str(cap) in msg
        # exactly the limit passes the size check (fails later, on the null pointers, with another error)
        rc = L.mgar_voxel_sa_fwd_static(cap // C1, 1, 1, C1, 1, 1, 0, None, None, None, None, None, None, None, 1, stream)
        assert rc == -1 and "null pointer" in L.mgar_last_error().decode()
