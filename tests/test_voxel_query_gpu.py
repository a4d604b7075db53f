"""-m gpu: voxel query through the C-ABI against the UNMODIFIED reference extension
(oracle/_ref/pointnet2_stack_cuda.so) - rows bit-exact - against the CPU oracle, and the modules built on it
(VoxelQueryAndGrouping, NeighborVoxelSAModuleMSG) against goldens produced by the reference class."""
import os

import numpy as np
import pytest
import torch

import cases
from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_stack import pointnet2_stack_cuda as new_pn2
from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_stack import voxel_pool_modules, voxel_query_utils
from multimodal_gar_b200.pcdet.utils import common_utils
from oracle import oracle as O

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def T(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _raw(mod, sc, dev, rng, radius, nsample):
    t = {k: T(sc[k], dev) for k in ("xyz", "new_xyz", "new_coords", "v2p")}
    M = sc["new_coords"].shape[0]
    _, Z, Y, X = sc["v2p"].shape
    idx = torch.zeros((M, nsample), dtype=torch.int32, device=dev)
    mod.voxel_query_wrapper(M, Z, Y, X, nsample, radius, rng[0], rng[1], rng[2], t["new_xyz"], t["xyz"], t["new_coords"],
                            t["v2p"], idx)
    return idx


CASES = [dict(seed=5, B=2, grid=(6, 24, 24), voxels=1500, queries_per_frame=90, rng=(2, 2, 2), radius=0.3, nsample=8),
         dict(seed=6, B=3, grid=(10, 60, 60), voxels=20000, queries_per_frame=4000, rng=(4, 4, 4), radius=0.45, nsample=16),
         dict(seed=7, B=1, grid=(3, 9, 200), voxels=3000, queries_per_frame=1001, rng=(0, 1, 5), radius=10.0, nsample=40),
         dict(seed=8, B=2, grid=(41, 100, 88), voxels=60000, queries_per_frame=27648, rng=(4, 4, 4), radius=0.8, nsample=16),
         dict(seed=9, B=2, grid=(5, 5, 5), voxels=50, queries_per_frame=33, rng=(7, 7, 7), radius=0.05, nsample=1)]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "seed%d" % c["seed"])
def test_voxel_query_bit_exact_vs_reference_extension_and_oracle(cuda, case):
    c = dict(case)
    rng, radius, nsample = c.pop("rng"), c.pop("radius"), c.pop("nsample")
    sc = cases.voxel_scene(C=1, **c)
    got = _raw(new_pn2, sc, cuda, rng, radius, nsample)
    orc = O.voxel_query_raw(rng, radius, nsample, sc["xyz"], sc["new_xyz"], sc["new_coords"], sc["v2p"])
    assert np.array_equal(got.cpu().numpy(), orc)
    ref = O.load_ref("pointnet2_stack_cuda")
    if ref is None:
        pytest.skip("oracle/_ref/pointnet2_stack_cuda.so not built")
    want = _raw(ref, sc, cuda, rng, radius, nsample)
    assert torch.equal(got, want)
    if case["seed"] != 9:
        empty = want[:, 0] == -1
        full = (want[:, -1] != want[:, 0]) & ~empty
        assert empty.any() and (~empty).any()
        assert full.any() or nsample >= 40


def test_voxel2pinds_and_grouping_module(cuda):
    sc = cases.voxel_scene(seed=11, B=2, grid=(6, 24, 24), voxels=1500, queries_per_frame=128, C=8)
    sp = type("Sp", (), dict(indices=T(sc["coords"], cuda), batch_size=2, spatial_shape=[6, 24, 24]))()
    v2p = common_utils.generate_voxel2pinds(sp)
    assert v2p.dtype == torch.int32 and np.array_equal(v2p.cpu().numpy(), sc["v2p"])
    g = voxel_query_utils.VoxelQueryAndGrouping((2, 2, 2), 0.3, 8)
    t = {k: T(v, cuda) for k, v in sc.items()}
    feat = t["features"].clone().requires_grad_(True)
    gf, gx, empty = g(t["new_coords"], t["xyz"], t["xyz_batch_cnt"], t["new_xyz"], t["new_xyz_batch_cnt"], feat, v2p)
    wf, wx, we = O.voxel_query_and_group((2, 2, 2), 0.3, 8, sc["new_coords"], sc["xyz"], sc["xyz_batch_cnt"], sc["new_xyz"],
                                         sc["new_xyz_batch_cnt"], sc["features"], sc["v2p"])
    assert np.array_equal(empty.cpu().numpy(), we) and we.any() and not we.all()
    assert np.array_equal(gf.detach().cpu().numpy(), wf) and np.array_equal(gx.cpu().numpy(), wx)
    w = torch.randn_like(gf)
    (gf * w).sum().backward()                                # gradient reaches the features through the grouping
    idx, emp = O.voxel_query((2, 2, 2), 0.3, 8, sc["xyz"], sc["new_xyz"], sc["new_coords"], sc["v2p"])
    start = np.concatenate([[0], np.cumsum(sc["xyz_batch_cnt"])[:-1]])
    idx[emp] = start[sc["new_coords"][emp, 0]][:, None]      # an empty row groups local index 0 of ITS frame (:89-91)
    want = np.zeros_like(sc["features"])
    np.add.at(want, idx.reshape(-1), w.permute(0, 2, 1).reshape(-1, 8).cpu().numpy())
    np.testing.assert_allclose(feat.grad.cpu().numpy(), want, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("pool", ["max_pool", "avg_pool"])
def test_neighbor_voxel_sa_module_vs_reference_class_golden(cuda, pool):
    g = np.load(os.path.join(GOLD, "ref_voxel_pool.npz"))
    m = voxel_pool_modules.NeighborVoxelSAModuleMSG(query_ranges=[[2, 2, 2], [1, 3, 3]], radii=[0.3, 0.45], nsamples=[8, 16],
                                                    mlps=[[16, 16, 24], [16, 32, 8]], pool_method=pool)
    sd = {k[len(pool) + 1:]: torch.from_numpy(g[k]) for k in g.files if k.startswith(pool + ".") and
          not k.endswith((".eval", ".train"))}
    m.load_state_dict(sd)                                     # the reference's parameter names, strict
    m = m.to(cuda)
    t = {k: T(g[k], cuda) for k in ("xyz", "xyz_batch_cnt", "new_xyz", "new_xyz_batch_cnt", "new_coords_xyz", "features", "v2p")}
    args = (t["xyz"], t["xyz_batch_cnt"], t["new_xyz"], t["new_xyz_batch_cnt"], t["new_coords_xyz"], t["features"], t["v2p"])
    tf32 = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False     # the golden is fp32 (CPU)
    try:
        _check_module(m, args, g, pool)
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = tf32
    with pytest.raises(NotImplementedError):
        m.pool_method = "sum_pool"
        m(*args)


def _check_module(m, args, g, pool):
    from multimodal_gar_b200 import _lib
    with torch.no_grad():
        m.eval()
        n0 = _lib.launch_count()
        out = m(*args)
        # eval mode: per scale the voxel query + the two fused kernels of csrc/voxel_sa.cu - no cuDNN convolution
        assert _lib.launch_count() - n0 == 3 * len(m.groupers)
        scale = float(np.abs(g[pool + ".eval"]).max())
        with torch.enable_grad():                            # the torch composition (autograd on) agrees with the fused path
            comp = m(*args)
        np.testing.assert_allclose(out.cpu().numpy(), comp.detach().cpu().numpy(), rtol=1e-5, atol=1e-5 * scale)
        np.testing.assert_allclose(out.cpu().numpy(), g[pool + ".eval"], rtol=1e-5, atol=1e-5 * scale)   # fp32 tolerance of north_star
        m.train()
        out = m(*args)
        np.testing.assert_allclose(out.cpu().numpy(), g[pool + ".train"], rtol=1e-4, atol=1e-4 * scale)
