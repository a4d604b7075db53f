"""-m gpu: training mode of the SA module through the C ABI (batch-statistics BatchNorm forward + the whole
backward, sa_mlp_train.cu) against the reference's own composition (QueryAndGroup -> Conv2d / BatchNorm2d / ReLU
-> pool under torch autograd, pointnet2_modules.py:78-112) on the same inputs: outputs, running statistics and
every gradient.  Tolerances: forward 1e-5 relative (north_star); gradients 1e-4 relative to the largest entry
(fp32 sums over ~10^4-10^5 positions in a different order than cuDNN/ATen)."""
import copy
import os

import numpy as np
import pytest
import torch

from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_stack import pointnet2_modules
import cases

pytestmark = pytest.mark.gpu


def T(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _close(got, want, rel):
    scale = max(1e-6, float(want.abs().max()))
    assert torch.allclose(got, want, rtol=rel, atol=rel * scale), float((got - want).abs().max()) / scale


def _run(mod, args, feats, w_out):
    f = feats.clone().requires_grad_(True)
    _, out = mod(args[0], args[1], args[2], args[3], f)
    (out * w_out).sum().backward()
    return out.detach(), f.grad.detach(), [p.grad.detach().clone() for p in mod.parameters()]


@pytest.mark.parametrize("radii,nsamples,mlps,pool,kernels", [
    ([0.6], [32], [[2, 64, 128]], "max_pool", "fast"),       # hot-path shape: sa_mlp_train_fast.cu, C0 = 5
    ([0.6], [32], [[1, 64, 128]], "max_pool", "fast"),       # the same with one feature channel (intensity), C0 = 4
    ([0.25], [32], [[1, 64, 128]], "max_pool", "fast"),      # small balls: duplicated samples, empty balls
    ([0.6], [32], [[2, 64, 128]], "max_pool", "tc"),         # layer-2 forward on tcgen05 (3xFP16) instead of CUDA cores
    ([0.25], [32], [[1, 64, 128]], "max_pool", "tc"),
    ([0.6], [32], [[2, 64, 128]], "max_pool", "generic"),    # the generic kernels on the same shape
    ([0.6], [32], [[2, 64, 128]], "avg_pool", "fast"),       # avg-pool is not specialised: falls through to generic
    ([0.3, 0.8], [8, 16], [[2, 16, 32], [2, 8, 24, 40]], "max_pool", "fast"),
    ([0.5], [12], [[2, 20]], "avg_pool", "fast"),
])
def test_training_forward_backward_match_torch_composition(cuda, monkeypatch, radii, nsamples, mlps, pool, kernels):
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    if kernels != "fast":
        monkeypatch.setenv("MGAR_SA_TRAIN_KERNELS", kernels)
    xyz, cnt, new_xyz, ncnt, feats = cases.ragged_case()
    feats = np.ascontiguousarray(feats[:, :mlps[0][0]])
    torch.manual_seed(11)
    native = pointnet2_modules.StackSAModuleMSG(radii=radii, nsamples=nsamples, mlps=copy.deepcopy(mlps),
                                                pool_method=pool).to(cuda)
    with torch.no_grad():
        for m in native.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.weight.normal_(1, 0.2)
                m.bias.normal_(0, 0.2)
    ref = copy.deepcopy(native)
    native.train(); ref.train()
    args = (T(xyz, cuda), T(cnt, cuda), T(new_xyz, cuda), T(ncnt, cuda))
    f = T(feats, cuda)
    w_out = torch.randn((new_xyz.shape[0], sum(m[-1] for m in mlps)), device=cuda)
    out_n, gf_n, gp_n = _run(native, args, f, w_out)
    os.environ["MGAR_SA_TRAIN"] = "torch"
    try:
        out_r, gf_r, gp_r = _run(ref, args, f, w_out)
    finally:
        del os.environ["MGAR_SA_TRAIN"]
    _close(out_n, out_r, 1e-5)
    _close(gf_n, gf_r, 1e-4)
    for a, b, (name, _) in zip(gp_n, gp_r, native.named_parameters()):
        _close(a, b, 1e-4)
    for (na, ba), (nb, bb) in zip(native.named_buffers(), ref.named_buffers()):
        if ba.dtype.is_floating_point:
            _close(ba, bb, 1e-5)
        else:
            assert torch.equal(ba, bb), na


def test_training_step_changes_weights_and_eval_uses_updated_statistics(cuda):
    """One SGD step through the native path, then eval(): the fused inference kernel must see the new running
    statistics and weights (the packed parameters are re-folded)."""
    xyz, cnt, new_xyz, ncnt, feats = cases.ragged_case()
    torch.manual_seed(5)
    mod = pointnet2_modules.StackSAModuleMSG(radii=[0.6], nsamples=[32], mlps=[[2, 64, 128]]).to(cuda)
    args = (T(xyz, cuda), T(cnt, cuda), T(new_xyz, cuda), T(ncnt, cuda), T(feats, cuda))
    mod.eval()
    with torch.no_grad():
        before = mod(*args)[1].clone()
    mod.train()
    opt = torch.optim.SGD(mod.parameters(), lr=0.1)
    loss = mod(*args)[1].square().mean()
    loss.backward()
    opt.step()
    mod.eval()
    with torch.no_grad():
        after = mod(*args)[1]
        os.environ["MGAR_SA_MLP"] = "fp32"
        try:
            mod._packed = None
            after32 = mod(*args)[1]
        finally:
            del os.environ["MGAR_SA_MLP"]
    assert not torch.allclose(before, after)
    _close(after, after32, 1e-5)


def test_fast_training_kernels_negative_gamma_and_uniform_batch(cuda):
    """Hot-path layout (uniform clouds of 512 points, 512 centroids each, one feature channel) with some NEGATIVE
    BatchNorm gammas in both layers: the specialised forward then takes the min over the samples for those
    channels (sa_mlp_train_fast.cu, identity 1).  6 persons -> 3072 centroids, 98 304 positions."""
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(3)
    Bp, S = 6, 512
    xyz = (torch.rand(Bp * S, 3, device=cuda) - 0.5) * torch.tensor([0.8, 0.8, 1.8], device=cuda)
    feats = torch.rand(Bp * S, 1, device=cuda)
    cnt = torch.full((Bp,), S, dtype=torch.int32, device=cuda)
    new_xyz = xyz.clone()
    native = pointnet2_modules.StackSAModuleMSG(radii=[0.3], nsamples=[32], mlps=[[1, 64, 128]]).to(cuda)
    with torch.no_grad():
        for m in native.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.weight.normal_(0.3, 1.0)
                m.bias.normal_(0, 0.3)
    ref = copy.deepcopy(native)
    native.train(); ref.train()
    args = (xyz, cnt, new_xyz, cnt)
    w_out = torch.randn((Bp * S, 128), device=cuda)
    out_n, gf_n, gp_n = _run(native, args, feats, w_out)
    os.environ["MGAR_SA_TRAIN"] = "torch"
    try:
        out_r, gf_r, gp_r = _run(ref, args, feats, w_out)
    finally:
        del os.environ["MGAR_SA_TRAIN"]
    _close(out_n, out_r, 1e-5)
    _close(gf_n, gf_r, 1e-4)
    for a, b in zip(gp_n, gp_r):
        _close(a, b, 1e-4)
    for (na, ba), (nb, bb) in zip(native.named_buffers(), ref.named_buffers()):
        if ba.dtype.is_floating_point:
            _close(ba, bb, 1e-5)
