"""-m gpu: the fused Distance Aware Fusion path (csrc/dafm.cu: mgar_dafm_attention + mgar_dafm_add_layernorm between the
library GEMMs) against the reference-class golden (tests/golden/ref_dafm.npz, 1e-5 of the largest feature) and against
the torch composition of the same module on the same GPU."""
import os

import numpy as np
import pytest
import torch

from multimodal_gar_b200 import _lib, dafm

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_dafm.npz"))


def _module(dev):
    D = int(G["D"])
    fus = dafm.DistanceAwareFusion(D)
    for name, mod in (("m1.", fus.AttFusModule1), ("m2.", fus.AttFusModule2)):
        mod.load_state_dict({k[len(name):]: torch.from_numpy(G[k]) for k in G.files if k.startswith(name)})
    return fus.to(dev).eval()


def test_fused_dafm_matches_reference_golden(cuda):
    fus = _module(cuda)
    counts = [int(c) for c in G["counts"]]
    MNP = G["rgb"].shape[1]
    valid = (torch.arange(MNP).view(1, -1) < torch.tensor(counts).view(-1, 1)).to(cuda)
    t = lambda k: torch.from_numpy(G[k]).to(cuda)
    n0 = _lib.launch_count()
    with torch.no_grad():
        got = fus(t("rgb"), t("lidar"), t("bboxes"), t("bboxes3d"), valid)
    assert _lib.launch_count() - n0 == 6, "two modules x (1 attention + 2 add-LayerNorm launches) expected"
    want = t("fused")
    scale = float(want.abs().max())
    assert torch.allclose(got, want, rtol=1e-5, atol=1e-5 * scale), float((got - want).abs().max())
    assert (got[~valid] == 0).all()


@pytest.mark.parametrize("B,n,D,full", [(3, 100, 512, False), (64, 100, 512, False), (2, 128, 512, True), (5, 7, 96, False), (1, 1, 512, True)])
def test_fused_module_matches_torch_composition(cuda, monkeypatch, B, n, D, full):
    torch.manual_seed(B * 1000 + n)
    mod = dafm.FusionAttentionMat(D, D, sigma=10).to(cuda).eval()
    with torch.no_grad():
        for p in mod.parameters():
            if p.dim() == 1:
                p.add_(0.1 * torch.randn_like(p))
    R, L = torch.randn(B, n, D, device=cuda), torch.randn(B, n, D, device=cuda)
    pos = torch.rand(B, n, 3, device=cuda) * 15
    De = torch.cdist(pos, pos)
    counts = torch.full((B,), n) if full else torch.randint(1, n + 1, (B,))
    valid = (torch.arange(n).view(1, -1) < counts.view(-1, 1)).to(cuda)
    tf32 = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        with torch.no_grad():
            mod(R, L, None, De, None if full else valid)        # (the first call packs the weights for the tcgen05 GEMM)
            n0 = _lib.launch_count()
            Rf, Lf = mod(R, L, None, De, None if full else valid)
            assert _lib.launch_count() - n0 == (6 if D % 128 == 0 else 3)   # + three grouped tcgen05 launches (six products) when the width fits
            monkeypatch.setenv("MGAR_DAFM", "torch")
            Rt, Lt = mod(R, L, None, De, None if full else valid)
            assert _lib.launch_count() - n0 == (6 if D % 128 == 0 else 3)   # + three grouped tcgen05 launches (six products) when the width fits
    finally:
        torch.backends.cuda.matmul.allow_tf32 = tf32
    for got, want in ((Rf, Rt), (Lf, Lt)):
        assert torch.allclose(got, want, rtol=1e-5, atol=2e-5 * float(want.abs().max())), float((got - want).abs().max())
        if not full:
            assert (got[~valid] == 0).all()


def test_autograd_keeps_the_torch_composition(cuda):
    mod = dafm.FusionAttentionMat(64, 64).to(cuda)
    R, L = torch.randn(2, 5, 64, device=cuda, requires_grad=True), torch.randn(2, 5, 64, device=cuda)
    De = torch.rand(2, 5, 5, device=cuda)
    n0 = _lib.launch_count()
    Rp, Lp = mod(R, L, None, De)
    (Rp.sum() + Lp.sum()).backward()
    assert _lib.launch_count() == n0 and R.grad is not None and torch.isfinite(R.grad).all()


def test_rejects_too_many_persons(cuda):
    mod = dafm.FusionAttentionMat(64, 64).to(cuda).eval()
    with torch.no_grad():
        R = torch.randn(1, 129, 64, device=cuda)
        n0 = _lib.launch_count()
        mod(R, R, None, torch.rand(1, 129, 129, device=cuda))                # falls to the torch composition (n > 128)
        assert _lib.launch_count() == n0
