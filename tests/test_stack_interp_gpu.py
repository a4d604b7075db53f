"""-m gpu: the stack-layout feature-propagation ops (three_nn / three_interpolate / three_interpolate_grad of
pointnet2_stack_cuda) through the C-ABI against the UNMODIFIED reference extension (oracle/_ref/pointnet2_stack_cuda.so)
- indices, squared distances and interpolated values bit-exact, the scatter gradient to float-atomic order - against
the CPU oracle, and StackPointnetFPModule against goldens produced by the reference class."""
import os

import numpy as np
import pytest
import torch

from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_stack import pointnet2_modules, pointnet2_utils
from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_stack import pointnet2_stack_cuda as new
from oracle import oracle as O

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def T(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _ref():
    mod = O.load_ref("pointnet2_stack_cuda")
    if mod is None:
        pytest.skip("oracle/_ref/pointnet2_stack_cuda.so not built")
    return mod


def _clouds(seed, ucnt, kcnt, dup=0):
    rng = np.random.default_rng(seed)
    ucnt, kcnt = np.asarray(ucnt, np.int32), np.asarray(kcnt, np.int32)
    unknown = (rng.random((int(ucnt.sum()), 3)) * np.array([4.0, 4.0, 1.5])).astype(np.float32)
    known = (rng.random((int(kcnt.sum()), 3)) * np.array([4.0, 4.0, 1.5])).astype(np.float32)
    if dup:
        known[rng.integers(0, kcnt[0], dup)] = known[rng.integers(0, kcnt[0], dup)]       # exact ties inside cloud 0
    return rng, unknown, ucnt, known, kcnt


CASES = [(0, (700, 333, 50), (128, 5, 77), 0),
         (1, (5000, 0, 129, 128, 1), (2049, 10, 3, 1024, 3), 40),        # empty unknown cloud, tile edges, ties
         (2, (100,), (100,), 100),
         (3, (40000, 25000), (4096, 9000), 0),
         (4, (17, 300), (2, 1), 0)]                                      # fewer than three known points: (+inf, cloud start)


@pytest.mark.parametrize("seed,ucnt,kcnt,dup", CASES)
def test_three_nn_stack(cuda, seed, ucnt, kcnt, dup):
    ref = _ref()
    _, unknown, ucnt, known, kcnt = _clouds(seed, ucnt, kcnt, dup)
    u, uc, k, kc = T(unknown, cuda), T(ucnt, cuda), T(known, cuda), T(kcnt, cuda)
    dw, dg = torch.zeros_like(u), torch.zeros_like(u)
    iw = torch.zeros(u.shape, dtype=torch.int32, device=cuda)
    ig = torch.zeros_like(iw)
    ref.three_nn_wrapper(u, uc, k, kc, dw, iw)
    new.three_nn_wrapper(u, uc, k, kc, dg, ig)
    assert torch.equal(ig, iw)
    assert torch.equal(dg, dw)
    if unknown.shape[0] * max(kcnt) <= 5e7:
        _, oi, od2 = O.stack_three_nn(unknown, ucnt, known, kcnt)
        assert np.array_equal(ig.cpu().numpy(), oi) and np.array_equal(dg.cpu().numpy(), od2)
    dist, idx = pointnet2_utils.three_nn(u, uc, k, kc)                    # the autograd Function returns sqrt
    assert torch.equal(idx, iw) and torch.equal(dist, torch.sqrt(dw))


@pytest.mark.parametrize("seed,N,M,C", [(0, 1083, 210, 7), (1, 20000, 5000, 128), (2, 1, 3, 1), (3, 4097, 64, 36), (4, 300, 50, 3)])
def test_three_interpolate_stack_and_gradient(cuda, seed, N, M, C):
    ref = _ref()
    rng = np.random.default_rng(seed)
    feats = rng.standard_normal((M, C)).astype(np.float32)
    idx = rng.integers(0, M, (N, 3)).astype(np.int32)
    weight = rng.random((N, 3)).astype(np.float32)
    f, i, w = T(feats, cuda), T(idx, cuda), T(weight, cuda)
    ow, og = torch.zeros((N, C), device=cuda), torch.full((N, C), 7.0, device=cuda)
    ref.three_interpolate_wrapper(f, i, w, ow)
    new.three_interpolate_wrapper(f, i, w, og)
    assert torch.equal(og, ow)
    assert np.array_equal(og.cpu().numpy(), O.stack_three_interpolate(feats, idx, weight))
    g = torch.randn_like(og)
    gw, gg = torch.zeros_like(f), torch.zeros_like(f)
    ref.three_interpolate_grad_wrapper(g, i, w, gw)
    new.three_interpolate_grad_wrapper(g, i, w, gg)
    torch.testing.assert_close(gg, gw, rtol=1e-4, atol=1e-4)
    np.testing.assert_allclose(gg.cpu().numpy(), O.stack_three_interpolate_grad(g.cpu().numpy(), idx, weight, M),
                               rtol=1e-4, atol=1e-4)
    f.requires_grad_(True)                                               # autograd Function: same scatter
    pointnet2_utils.three_interpolate(f, i, w).backward(g)
    torch.testing.assert_close(f.grad, gw, rtol=1e-4, atol=1e-4)


def test_unaligned_feature_rows_take_the_scalar_path(cuda):
    rng = np.random.default_rng(5)
    M, N, C = 40, 257, 8
    buf = T(rng.standard_normal(M * C + 1).astype(np.float32), cuda)
    f = buf[1:].view(M, C)                                               # C % 4 == 0 but the base is 4-byte aligned only
    idx = rng.integers(0, M, (N, 3)).astype(np.int32)
    weight = rng.random((N, 3)).astype(np.float32)
    out = torch.zeros((N, C), device=cuda)
    new.three_interpolate_wrapper(f, T(idx, cuda), T(weight, cuda), out)
    assert np.array_equal(out.cpu().numpy(), O.stack_three_interpolate(f.cpu().numpy(), idx, weight))


def test_fp_module_vs_reference_class_goldens(cuda):
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_ref_stack_fp_golden", os.path.join(GOLD, "make_ref_stack_fp_golden.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    inputs = gen.inputs
    g = np.load(os.path.join(GOLD, "ref_stack_fp.npz"))
    unknown, ucnt, known, kcnt, ufeat, kfeat = inputs()
    fp = pointnet2_modules.StackPointnetFPModule(mlp=[kfeat.shape[1] + ufeat.shape[1], 32, 16])
    fp_nofeat = pointnet2_modules.StackPointnetFPModule(mlp=[kfeat.shape[1], 8])
    for name, m in (("fp", fp), ("fp_nofeat", fp_nofeat)):
        m.load_state_dict({k[len(name) + 1:]: torch.from_numpy(g[k]) for k in g.files if k.startswith(name + ".")})
        m.to(cuda)
    u, uc, k, kc, uf, kf = (T(a, cuda) for a in (unknown, ucnt, known, kcnt, ufeat, kfeat))
    tf32 = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False     # the goldens are fp32 (CPU)
    try:
        with torch.no_grad():
            for mode, tol in (("eval", 1e-5), ("train", 1e-4)):
                fp.train(mode == "train")
                fp_nofeat.train(mode == "train")
                for got, key in ((fp(u, uc, k, kc, uf, kf), "fp"), (fp_nofeat(u, uc, k, kc, None, kf), "fp_nofeat")):
                    want = g["%s.%s" % (mode, key)]
                    np.testing.assert_allclose(got.cpu().numpy(), want, rtol=tol, atol=tol * float(np.abs(want).max()))
        fp.eval()
        kf.requires_grad_(True)
        fp(u, uc, k, kc, uf, kf).square().sum().backward()
        assert kf.grad is not None and float(kf.grad.abs().sum()) > 0 and torch.isfinite(kf.grad).all()
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = tf32
