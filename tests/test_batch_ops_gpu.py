"""-m gpu: the pointnet2_batch surface through the C-ABI against the UNMODIFIED reference extension
(oracle/_ref/pointnet2_batch_cuda.so) - indices and gathered / interpolated values bit-exact, scatter gradients to
float-atomic order - against the CPU oracle, and the modules against goldens produced by the reference classes."""
import os

import numpy as np
import pytest
import torch

from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_batch import pointnet2_batch_cuda as new
from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_batch import pointnet2_modules, pointnet2_utils
from oracle import oracle as O

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def T(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _ref():
    mod = O.load_ref("pointnet2_batch_cuda")
    if mod is None:
        pytest.skip("oracle/_ref/pointnet2_batch_cuda.so not built")
    return mod


def _cloud(seed, B, N, scale=(4.0, 4.0, 1.5), dup=0):
    rng = np.random.default_rng(seed)
    xyz = (rng.random((B, N, 3)) * np.asarray(scale)).astype(np.float32)
    if dup:
        xyz[:, rng.integers(0, N, dup)] = xyz[:, rng.integers(0, N, dup)]        # exact ties
    return rng, xyz


@pytest.mark.parametrize("seed,B,N,M,radius,ns", [(0, 2, 700, 96, 0.3, 8), (1, 3, 16384, 4096, 0.2, 32), (2, 1, 511, 511, 5.0, 64),
                                                  (3, 4, 33, 7, 0.01, 5), (4, 2, 100000, 512, 0.1, 16)])
def test_ball_query_batch(cuda, seed, B, N, M, radius, ns):
    ref = _ref()
    rng, xyz = _cloud(seed, B, N)
    new_xyz = np.ascontiguousarray(xyz[:, rng.permutation(N)[:M]])
    new_xyz[:, :2] += 50.0                                                       # centroids without neighbours
    x, q = T(xyz, cuda), T(new_xyz, cuda)
    want = torch.zeros((B, M, ns), dtype=torch.int32, device=cuda)
    got = torch.zeros_like(want)
    ref.ball_query_wrapper(B, N, M, radius, ns, q, x, want)
    new.ball_query_wrapper(B, N, M, radius, ns, q, x, got)
    assert torch.equal(got, want)
    assert (want[:, :2] == 0).all() and (want >= 0).all()                        # no -1 marker in this module
    if N <= 20000:
        assert np.array_equal(got.cpu().numpy(), O.batch_ball_query(radius, ns, xyz, new_xyz))


@pytest.mark.parametrize("seed,B,C,N,P,ns", [(0, 2, 5, 700, 96, 8), (1, 3, 67, 16384, 1024, 32), (2, 1, 1, 50, 1, 1), (3, 2, 9, 300, 77, 1)])
def test_group_and_gather_points_and_gradients(cuda, seed, B, C, N, P, ns):
    ref = _ref()
    rng = np.random.default_rng(seed)
    feats = rng.standard_normal((B, C, N)).astype(np.float32)
    idx = rng.integers(0, N, (B, P, ns)).astype(np.int32)
    idx[:, 0] = idx[:, 0, :1]                                                    # repeated indices (padded rows)
    f, i = T(feats, cuda), T(idx, cuda)
    want, got = torch.empty((B, C, P, ns), device=cuda), torch.empty((B, C, P, ns), device=cuda)
    ref.group_points_wrapper(B, C, N, P, ns, f, i, want)
    new.group_points_wrapper(B, C, N, P, ns, f, i, got)
    assert torch.equal(got, want)
    assert np.array_equal(got.cpu().numpy(), O.batch_grouping_operation(feats, idx))
    g = torch.randn_like(got)
    gw, gg = torch.zeros_like(f), torch.zeros_like(f)
    ref.group_points_grad_wrapper(B, C, N, P, ns, g, i, gw)
    new.group_points_grad_wrapper(B, C, N, P, ns, g, i, gg)
    torch.testing.assert_close(gg, gw, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(gg.cpu().numpy(), O.batch_grouping_grad(g.cpu().numpy(), idx, N), rtol=1e-5, atol=1e-5)
    if ns == 1:                                                                  # gather_points: the same op, one sample
        i1 = i.view(B, P).contiguous()
        w1, g1 = torch.empty((B, C, P), device=cuda), torch.empty((B, C, P), device=cuda)
        ref.gather_points_wrapper(B, C, N, P, f, i1, w1)
        new.gather_points_wrapper(B, C, N, P, f, i1, g1)
        assert torch.equal(g1, w1)
        g = torch.randn_like(g1)
        gw, gg = torch.zeros_like(f), torch.zeros_like(f)
        ref.gather_points_grad_wrapper(B, C, N, P, g, i1, gw)
        new.gather_points_grad_wrapper(B, C, N, P, g, i1, gg)
        torch.testing.assert_close(gg, gw, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("seed,B,N,M", [(0, 2, 700, 96), (1, 2, 16384, 4096), (2, 1, 5, 2), (3, 3, 1500, 1025)])
def test_fps_batch_entry(cuda, seed, B, N, M):
    ref = _ref()
    _, xyz = _cloud(seed, B, N, dup=N // 10)
    x = T(xyz, cuda)
    want, got = torch.zeros((B, M), dtype=torch.int32, device=cuda), torch.zeros((B, M), dtype=torch.int32, device=cuda)
    ref.farthest_point_sampling_wrapper(B, N, M, x, torch.full((B, N), 1e10, device=cuda), want)
    new.farthest_point_sampling_wrapper(B, N, M, x, torch.full((B, N), 1e10, device=cuda), got)
    assert torch.equal(got, want)


@pytest.mark.parametrize("seed,B,N,M,C", [(0, 2, 700, 96, 56), (1, 2, 16384, 4096, 128), (2, 1, 40, 2, 3), (3, 3, 2000, 1025, 9), (4, 1, 9, 1, 2)])
def test_three_nn_and_interpolate(cuda, seed, B, N, M, C):
    ref = _ref()
    rng, unknown = _cloud(seed, B, N)
    known = np.ascontiguousarray(unknown[:, rng.permutation(N)[:M]])             # exact zero distances
    if M > 8:
        known[:, 5] = known[:, 3]                                                # tie between two known points
    u, k = T(unknown, cuda), T(known, cuda)
    dw, dg = torch.zeros((B, N, 3), device=cuda), torch.zeros((B, N, 3), device=cuda)
    iw, ig = torch.zeros((B, N, 3), dtype=torch.int32, device=cuda), torch.zeros((B, N, 3), dtype=torch.int32, device=cuda)
    ref.three_nn_wrapper(B, N, M, u, k, dw, iw)
    new.three_nn_wrapper(B, N, M, u, k, dg, ig)
    assert torch.equal(ig, iw) and torch.equal(dg, dw)
    _, oi, od2 = O.three_nn(unknown, known)
    assert np.array_equal(ig.cpu().numpy(), oi) and np.array_equal(dg.cpu().numpy(), od2)
    if M < 3:
        assert torch.isinf(dg[..., M:]).all() and (ig[..., M:] == 0).all()
    feats = rng.standard_normal((B, C, M)).astype(np.float32)
    weight = rng.random((B, N, 3)).astype(np.float32)
    f, w = T(feats, cuda), T(weight, cuda)
    ow, og = torch.empty((B, C, N), device=cuda), torch.empty((B, C, N), device=cuda)
    ref.three_interpolate_wrapper(B, C, M, N, f, iw, w, ow)
    new.three_interpolate_wrapper(B, C, M, N, f, iw, w, og)
    assert torch.equal(og, ow)
    assert np.array_equal(og.cpu().numpy(), O.three_interpolate(feats, oi, weight))
    g = torch.randn_like(og)
    gw, gg = torch.zeros_like(f), torch.zeros_like(f)
    ref.three_interpolate_grad_wrapper(B, C, N, M, g, iw, w, gw)
    new.three_interpolate_grad_wrapper(B, C, N, M, g, iw, w, gg)
    torch.testing.assert_close(gg, gw, rtol=1e-4, atol=1e-4)
    np.testing.assert_allclose(gg.cpu().numpy(), O.three_interpolate_grad(g.cpu().numpy(), oi, weight, M), rtol=1e-4, atol=1e-4)


def test_modules_vs_reference_class_goldens_and_autograd(cuda):
    g = np.load(os.path.join(GOLD, "ref_batch_modules.npz"))
    C = g["feats"].shape[1]
    sa = pointnet2_modules.PointnetSAModuleMSG(npoint=96, radii=[0.3, 0.6], nsamples=[8, 16], mlps=[[C, 16, 24], [C, 16, 32]])
    sa_all = pointnet2_modules.PointnetSAModule(mlp=[C, 32], use_xyz=True, pool_method="avg_pool")
    fp = pointnet2_modules.PointnetFPModule(mlp=[56 + C, 32, 16])
    for name, m in (("sa", sa), ("sa_all", sa_all), ("fp", fp)):
        m.load_state_dict({k[len(name) + 1:]: torch.from_numpy(g[k]) for k in g.files if k.startswith(name + ".")})
        m.to(cuda)
    xyz, feats = T(g["xyz"], cuda), T(g["feats"], cuda)
    tf32 = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False     # the goldens are fp32 (CPU)
    try:
        with torch.no_grad():
            for mode, tol in (("eval", 1e-5), ("train", 1e-4)):
                for m in (sa, sa_all, fp):
                    m.train(mode == "train")
                new_xyz, f1 = sa(xyz, feats)
                _, f2 = sa_all(xyz, feats)
                up = fp(xyz, new_xyz, feats, f1)
                assert np.array_equal(new_xyz.cpu().numpy(), g[mode + ".new_xyz"])
                for got, key in ((f1, "sa"), (f2, "sa_all"), (up, "fp")):
                    want = g["%s.%s" % (mode, key)]
                    np.testing.assert_allclose(got.cpu().numpy(), want, rtol=tol, atol=tol * float(np.abs(want).max()))
        feats.requires_grad_(True)                               # gradients reach the features through every op
        for m in (sa, fp):
            m.eval()
        new_xyz, f1 = sa(xyz, feats)
        up = fp(xyz, new_xyz, feats, f1)
        up.square().sum().backward()
        assert feats.grad is not None and float(feats.grad.abs().sum()) > 0 and torch.isfinite(feats.grad).all()
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = tf32


def test_utils_functions_match_oracle(cuda):
    rng, xyz = _cloud(9, 2, 600)
    x = T(xyz, cuda)
    idx = pointnet2_utils.farthest_point_sample(x, 64)
    assert np.array_equal(idx.cpu().numpy(), O.farthest_point_sample(xyz, 64))
    new_xyz = pointnet2_utils.gather_operation(x.transpose(1, 2).contiguous(), idx).transpose(1, 2).contiguous()
    bq = pointnet2_utils.ball_query(0.4, 12, x, new_xyz)
    assert np.array_equal(bq.cpu().numpy(), O.batch_ball_query(0.4, 12, xyz, new_xyz.cpu().numpy()))
    feats = T(rng.standard_normal((2, 6, 600)).astype(np.float32), cuda)
    grouped = pointnet2_utils.QueryAndGroup(0.4, 12)(x, new_xyz, feats)
    assert grouped.shape == (2, 9, 64, 12)
    assert pointnet2_utils.GroupAll()(x, None, feats).shape == (2, 9, 1, 600)
    with pytest.raises(Exception):
        new.three_nn_wrapper(2, 600, 64, x.cpu(), new_xyz, torch.zeros(2, 600, 3), torch.zeros(2, 600, 3, dtype=torch.int32))
