"""-m gpu: the 3xTF32 tcgen05 GEMM (csrc/gemm_tc.cu) against an fp64 product of the same fp32 operands: 1e-5 relative to
the row's scale (north_star's bar for fp32 features; a single-pass TF32 GEMM misses it by two orders of magnitude, which
the test also shows), bias / ReLU epilogue, ragged M, strided operands, and the DAFM / embedding call sites."""
import numpy as np
import pytest
import torch

from multimodal_gar_b200 import _lib, dafm, tc_gemm

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("M,N,K,bias,relu", [(6400, 1536, 512, False, False), (6400, 512, 512, True, True), (1, 128, 32, True, False),
                                             (129, 256, 64, False, True), (5000, 512, 128, True, False), (300, 128, 2048, True, True)])
def test_gemm_matches_fp64_product(cuda, M, N, K, bias, relu):
    g = torch.Generator(device="cpu").manual_seed(M + N + K)
    x = (torch.randn((M, K), generator=g) * torch.logspace(-3, 3, K)).to(cuda)        # wide dynamic range along K
    w = torch.randn((N, K), generator=g).to(cuda) / K ** 0.5
    b = torch.randn((N,), generator=g).to(cuda) if bias else None
    y0 = tc_gemm.linear(x, w, b, relu=relu, packed=False)
    tc_gemm.linear(x, w, b, relu=relu)                     # (packs the weight on first use)
    n0 = _lib.launch_count()
    y = tc_gemm.linear(x, w, b, relu=relu)
    assert _lib.launch_count() == n0 + 1                   # packed weights are cached: one launch per call
    assert torch.equal(y, y0)                              # same split, same accumulation order: same bits
    want = x.double() @ w.double().t()
    if bias:
        want = want + b.double()
    if relu:
        want = want.clamp_min(0)
    mag = (x.double().abs() @ w.double().abs().t()).amax(dim=1, keepdim=True)          # scale of a row's products
    err = ((y.double() - want).abs() / mag).max().item()
    assert err < 4e-6, err
    torch.backends.cuda.matmul.allow_tf32 = False
    ref32 = torch.nn.functional.linear(x, w, b)
    ref32 = ref32.clamp_min(0) if relu else ref32
    err32 = ((ref32.double() - want).abs() / mag).max().item()
    assert err < 8 * max(err32, 1e-7), (err, err32)                                    # as good as the fp32 library GEMM


def test_gemm_strided_rows_and_rejects_bad_shapes(cuda):
    x = torch.randn((70, 96), device=cuda)[:, :64]                                     # row stride 96
    w = torch.randn((128, 64), device=cuda)
    def close(y, x, w):                                    # against the fp64 product, relative to a row's |x|.|w|
        mag = (x.double().abs() @ w.double().abs().t()).amax(dim=1, keepdim=True)
        return float(((y.double() - x.double() @ w.double().t()).abs() / mag).max()) < 4e-6
    assert close(tc_gemm.linear(x, w), x, w)
    assert close(tc_gemm.linear(x, w, packed=False), x, w)
    w.mul_(2.0)                                            # in-place edit: the packed copy is keyed on the version
    assert close(tc_gemm.linear(x, w), x, w)
    assert not tc_gemm.usable(x, 100, 64) and not tc_gemm.usable(x, 128, 40)
    with pytest.raises(_lib.MgarError):
        tc_gemm.linear(torch.randn((8, 40), device=cuda), torch.randn((128, 40), device=cuda))


def test_dafm_runs_its_linears_on_the_tensor_cores(cuda, monkeypatch):
    """The fused DAFM path runs its 6 products per module (2 projections + 4 FFN linears) as 3 grouped tcgen05 launches (the
    two streams of a layer share a launch) and no library GEMM, and agrees with the torch composition at 1e-5."""
    torch.manual_seed(2)
    B, n, D = 5, 37, 512
    mod = dafm.FusionAttentionMat(D, D).to(cuda).eval()
    R, L = torch.randn((B, n, D), device=cuda), torch.randn((B, n, D), device=cuda)
    ctr = torch.randn((B, n, 3), device=cuda) * 5
    De = torch.cdist(ctr, ctr)
    valid = torch.arange(n, device=cuda).view(1, n) < torch.tensor([37, 20, 1, 36, 5], device=cuda).view(B, 1)
    with torch.no_grad():
        mod(R, L, None, De, valid)                         # first call packs the weights
        n0 = _lib.launch_count()
        got = mod(R, L, None, De, valid)
        launches = _lib.launch_count() - n0
        monkeypatch.setenv("MGAR_DAFM", "torch")
        torch.backends.cuda.matmul.allow_tf32 = False
        want = mod(R, L, None, De, valid)
    assert launches == 3 + 1 + 2, launches                                             # grouped GEMMs + attention + 2 add-LayerNorm
    for a, b in zip(got, want):
        scale = float(b.abs().max())
        assert torch.allclose(a, b, rtol=1e-5, atol=2e-5 * scale), float((a - b).abs().max()) / scale
    # a parameter edited in place is picked up (the K-major weight copies are keyed on the parameter versions)
    with torch.no_grad():
        mod.WQ_r.mul_(1.5)
        monkeypatch.delenv("MGAR_DAFM")
        again = mod(R, L, None, De, valid)
        monkeypatch.setenv("MGAR_DAFM", "torch")
        want2 = mod(R, L, None, De, valid)
    assert not torch.equal(again[0], got[0])
    assert torch.allclose(again[0], want2[0], rtol=1e-5, atol=2e-5 * float(want2[0].abs().max()))


def test_grouped_launch_equals_the_single_products(cuda):
    """mgar_gemm_tf32x3_packed_grouped: up to four products of one shape in one launch - the same bits as one launch each;
    mixed shapes fall back to single launches."""
    from multimodal_gar_b200 import tc_gemm
    g = torch.Generator(device="cpu").manual_seed(4)
    M, K, N = 777, 512, 512
    items = []
    for i in range(4):
        x = torch.randn((M, K), generator=g).to(cuda)
        w = torch.randn((N, K), generator=g).to(cuda)
        b = None if i == 2 else torch.randn((N,), generator=g).to(cuda)
        items.append((x, w, b))
        tc_gemm.packed_weight(w)                           # (the first use of a weight packs it: one more launch)
    for relu in (False, True):
        for cnt in (2, 3, 4):
            n0 = _lib.launch_count()
            got = tc_gemm.linear_grouped(items[:cnt], relu=relu)
            assert _lib.launch_count() - n0 == 1
            for (x, w, b), y in zip(items[:cnt], got):
                assert torch.equal(y, tc_gemm.linear(x, w, b, relu=relu))
    mixed = items[:1] + [(items[1][0][:100], items[1][1], items[1][2])]
    got = tc_gemm.linear_grouped(mixed)
    assert torch.equal(got[1], tc_gemm.linear(*mixed[1]))
