"""Distance Aware Fusion Module, batched (the step right after the LiDAR hot path).

Reference: `FusionAttention_mat` (/root/reference/model/gat_model.py:427-511) applied per frame inside a Python
loop by `GAR_Fusion_Net3.forward` (:1396-1475, FUSION == 'Attention_mat': two modules, then an element-wise max
of the RGB and LiDAR streams).  Same parameter names and shapes, so reference checkpoints load unchanged.

Here every frame of the batch goes through at once: the six 512x512 projections and the four FFN linears of a
module are single `(B*n, 512) x (512, 512)` library GEMMs, the n x n attention pieces are batched, and frames with
fewer persons than the padded width are handled by masking keys with -inf before each softmax, so that
`out[b, :n_b]` is what the reference loop computes for frame b (tests/test_dafm.py pins this against the
reference class itself).  On CUDA tensors without autograd everything between the GEMMs runs in libmgar_lidar.so
(csrc/dafm.cu): the projections become two (B*n, 512) x (512, 1536) GEMMs (the three weights sharing a left operand
side by side), one launch per module does scores * E / sqrt(D), key mask, softmax, Att V + residual and LayerNorm for
both streams (the n x n map stays in shared memory), and LayerNorm(x + FFN(x)) with the padded rows zeroed is one more
kernel per stream.  The torch composition below stays for CPU tensors (the CPU golden test) and for autograd.  Note two reference quirks that are kept: both streams use E_r = softmax(-De/sigma)
(Dg is accepted but unused, `E_l = Dg` is dead code :499), and softmax runs over keys (dim=1 of the (n, n) map).
"""
import ctypes
import os

import torch
import torch.nn as nn


def generalized_box_iou(boxes):
    """torchvision.ops.generalized_box_iou(b, b) batched: boxes (..., n, 4) as (x1, y1, x2, y2) -> (..., n, n)."""
    a, b = boxes.unsqueeze(-2), boxes.unsqueeze(-3)
    area = (boxes[..., 2] - boxes[..., 0]) * (boxes[..., 3] - boxes[..., 1])
    lt = torch.max(a[..., :2], b[..., :2])
    rb = torch.min(a[..., 2:], b[..., 2:])
    wh = (rb - lt).clamp(min=0)
    inter = wh[..., 0] * wh[..., 1]
    union = area.unsqueeze(-1) + area.unsqueeze(-2) - inter
    iou = inter / union
    lti = torch.min(a[..., :2], b[..., :2])
    rbi = torch.max(a[..., 2:], b[..., 2:])
    whi = (rbi - lti).clamp(min=0)
    areai = whi[..., 0] * whi[..., 1]
    return iou - (areai - union) / areai


def _masked_softmax(x, key_valid):
    """softmax over the last dim with invalid keys removed; x (B,n,n), key_valid (B,n) bool or None."""
    if key_valid is not None:
        x = x.masked_fill(~key_valid.unsqueeze(1), float("-inf"))
    return torch.softmax(x, dim=-1)


class FusionAttentionMat(nn.Module):
    """`FusionAttention_mat` (gat_model.py:427-511).  forward(R, L, Dg, De[, valid]) -> (R', L').

    Unbatched (n, D) inputs reproduce the reference call exactly; batched (B, n, D) inputs take `valid` (B, n) bool
    (True = real person); rows of padded persons come out as zeros."""

    def __init__(self, input_dim=512, out_dim=512, sigma=10):
        super().__init__()
        self.input_dim, self.out_dim, self.sigma = input_dim, out_dim, sigma
        shape = [input_dim, out_dim]
        for s in ("r", "l"):
            for n in ("WQ", "WK", "WV"):
                p = nn.Parameter(torch.zeros(shape))
                torch.nn.init.kaiming_normal_(p)
                setattr(self, "%s_%s" % (n, s), p)
            setattr(self, "LN_%s_1" % s, nn.LayerNorm([out_dim]))
            setattr(self, "FFN_%s" % s, nn.Sequential(nn.Linear(out_dim, out_dim), nn.ReLU(), nn.Linear(out_dim, out_dim)))
            setattr(self, "LN_%s_2" % s, nn.LayerNorm([out_dim]))
        from . import tc_gemm
        self._wl, self._wr = tc_gemm.CachedTranspose(), tc_gemm.CachedTranspose()   # K-major copies of the projections

    def _stream(self, q_src, kv_src, E, valid, s):
        WQ, WK, WV = getattr(self, "WQ_" + s), getattr(self, "WK_" + s), getattr(self, "WV_" + s)
        Q, K, V = q_src @ WQ, kv_src @ WK, kv_src @ WV
        att = _masked_softmax((Q @ K.transpose(-1, -2)) * E / self.out_dim ** 0.5, valid)
        x = getattr(self, "LN_%s_1" % s)(att @ V + kv_src)
        x = x + getattr(self, "FFN_" + s)(x)
        return getattr(self, "LN_%s_2" % s)(x)

    def _fused_ok(self, R, L):
        if not (R.is_cuda and L.is_cuda and R.dtype == L.dtype == torch.float32):
            return False
        if os.environ.get("MGAR_DAFM", "fused") == "torch" or self.input_dim != self.out_dim:
            return False
        if torch.is_grad_enabled() and (R.requires_grad or L.requires_grad or self.WQ_r.requires_grad):
            return False
        return R.shape[1] <= 128 and self.out_dim <= 512

    def _fused(self, R, L, E, valid):
        """Both streams through libmgar_lidar.so: 2 projection GEMMs, mgar_dafm_attention, FFN GEMMs,
        mgar_dafm_add_layernorm per stream."""
        from . import _lib, tc_gemm
        from ._torch_glue import on_device, ptr, stream_of
        B, n, D = R.shape
        R, L, E = R.contiguous(), L.contiguous(), E.contiguous()
        # the six projections as two (B*n, D) x (D, 3D) products, the four FFN linears as they are: on tcgen05 (3xTF32,
        # csrc/gemm_tc.cu) when the widths fit the kernel, else torch's fp32 GEMM (MGAR_NO_TC_GEMM=1 forces that)
        tc = tc_gemm.usable(L, 3 * D, D)
        if tc:
            # (the products of the two streams have one shape: one grouped launch each, their tiles fill the SMs together)
            PL, PR = tc_gemm.linear_grouped([(L.view(B * n, D), self._wl.get((self.WQ_r, self.WK_l, self.WV_l)), None),    # Q_r | K_l | V_l
                                             (R.view(B * n, D), self._wr.get((self.WK_r, self.WV_r, self.WQ_l)), None)])   # K_r | V_r | Q_l
        else:
            PL = L.view(B * n, D) @ torch.cat([self.WQ_r, self.WK_l, self.WV_l], dim=1)
            PR = R.view(B * n, D) @ torch.cat([self.WK_r, self.WV_r, self.WQ_l], dim=1)
        X = torch.empty((2, B, n, D), dtype=torch.float32, device=R.device)
        vmask = None if valid is None else valid.contiguous()
        pl, pr, f4 = PL.data_ptr(), PR.data_ptr(), 4 * D

        def arr(a, b):
            return (ctypes.c_void_p * 2)(a, b)

        L_ = _lib.lib()
        with on_device(R):
            st = stream_of(R)
            _lib.check(L_.mgar_dafm_attention(B, n, D, 2, 3 * D, arr(pl, pr + 2 * f4), arr(pr, pl + f4), arr(pr + f4, pl + 2 * f4),
                                              arr(ptr(R), ptr(L)), ptr(E), ptr(vmask),
                                              arr(ptr(self.LN_r_1.weight), ptr(self.LN_l_1.weight)),
                                              arr(ptr(self.LN_r_1.bias), ptr(self.LN_l_1.bias)), float(self.LN_r_1.eps),
                                              arr(X[0].data_ptr(), X[1].data_ptr()), st), "dafm_attention")
            outs = []
            ys = None
            if tc:
                xs = [X[i].view(B * n, D) for i in range(2)]
                ffns = [self.FFN_r, self.FFN_l]
                hs = tc_gemm.linear_grouped([(x, f[0].weight, f[0].bias) for x, f in zip(xs, ffns)], relu=True)
                ys = tc_gemm.linear_grouped([(h, f[2].weight, f[2].bias) for h, f in zip(hs, ffns)])
            for i, s in enumerate(("r", "l")):
                x = X[i].view(B * n, D)
                ffn = getattr(self, "FFN_" + s)
                y = ys[i] if tc else ffn(x)
                ln = getattr(self, "LN_%s_2" % s)
                o = torch.empty_like(x)
                _lib.check(L_.mgar_dafm_add_layernorm(B * n, D, ptr(x), ptr(y), ptr(ln.weight), ptr(ln.bias), ptr(vmask),
                                                      float(ln.eps), ptr(o), st), "dafm_add_layernorm")
                outs.append(o.view(B, n, D))
        return outs

    def distance_weights(self, De, valid=None):
        """E_r = softmax(-(De / sigma), dim=1) (:484) over the valid keys; De (B,n,n)."""
        return _masked_softmax(-(De / self.sigma), valid)

    def forward(self, R, L, Dg, De, valid=None, E=None):
        """E: optional precomputed `distance_weights(De, valid)` (batched call only) - the two modules of the fusion
        step share it."""
        squeeze = R.dim() == 2
        if squeeze:
            R, L, De = R.unsqueeze(0), L.unsqueeze(0), De.unsqueeze(0)
            valid = None if valid is None else valid.unsqueeze(0)
            E = None
        if E is None:
            E = self.distance_weights(De, valid)
        if self._fused_ok(R, L):
            Rp, Lp = self._fused(R, L, E, valid)
            return (Rp[0], Lp[0]) if squeeze else (Rp, Lp)
        Rp = self._stream(L, R, E, valid, "r")          # queries from LiDAR, keys / values from RGB (:480-482)
        Lp = self._stream(R, L, E, valid, "l")          # queries from RGB, keys / values from LiDAR (:495-497)
        if valid is not None:
            keep = valid.unsqueeze(-1)
            Rp, Lp = torch.where(keep, Rp, torch.zeros_like(Rp)), torch.where(keep, Lp, torch.zeros_like(Lp))
        return (Rp[0], Lp[0]) if squeeze else (Rp, Lp)


class DistanceAwareFusion(nn.Module):
    """FUSION == 'Attention_mat' of `GAR_Fusion_Net3.forward` (gat_model.py:1467-1475), all frames at once:
    (RGB_feature, LiDAR_feature (B,MNP,D), bboxes (B,MNP,4), bboxes3d (B,MNP,7), valid (B,MNP)) -> fused (B,MNP,D),
    zeros in padded rows.  Module names follow the reference (`AttFusModule1`, `AttFusModule2`)."""

    def __init__(self, dim=512, sigma=10):
        super().__init__()
        self.AttFusModule1 = FusionAttentionMat(dim, dim, sigma)
        self.AttFusModule2 = FusionAttentionMat(dim, dim, sigma)

    def forward(self, rgb, lidar, bboxes, bboxes3d, valid):
        from .lidar_backbone import pairwise_euclidean_distance
        # The reference also builds Dg = generalized_box_iou(bboxes) here (:1470) and hands it to both modules, which
        # never read it (`E_l = Dg` is dead code, :499): ~15 elementwise launches on (B,n,n) that cannot change the result.
        Dg = None
        De = pairwise_euclidean_distance(bboxes3d[..., :3])
        E1 = self.AttFusModule1.distance_weights(De, valid)
        E2 = E1 if self.AttFusModule2.sigma == self.AttFusModule1.sigma else None   # same De, same sigma: same weights
        R, L = self.AttFusModule1(rgb, lidar, Dg, De, valid, E=E1)
        R, L = self.AttFusModule2(R, L, Dg, De, valid, E=E2)
        return torch.maximum(R, L)
