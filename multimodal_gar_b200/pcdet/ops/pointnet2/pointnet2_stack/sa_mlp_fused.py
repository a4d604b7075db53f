"""Fused QueryAndGroup + shared MLP + pool (eval-mode BatchNorm) over libmgar_lidar.so.

This replaces, for inference, the chain in the reference's StackSAModuleMSG.forward
(/root/reference/pcdet/ops/pointnet2/pointnet2_stack/pointnet2_modules.py:89-110): grouping,
`cat`, [Conv2d 1x1, BatchNorm2d, ReLU] x L, max_pool2d.
"""
import ctypes
import os

import torch
import torch.nn as nn

from ..... import _lib
from ....._torch_glue import need_cuda, need_dtype, on_device, ptr, stream_of


def pack_shared_mlp(seq):
    """nn.Sequential([Conv2d(1x1,no bias), BatchNorm2d, ReLU] x L) -> (channels list, packed fp32 tensor).

    Per layer: W^T [c_in][c_out], scale[c_out], shift[c_out] with the running statistics folded:
    scale = gamma / sqrt(var + eps), shift = beta - mean * scale."""
    mods = list(seq)
    assert len(mods) % 3 == 0 and len(mods) > 0
    channels, parts = [], []
    for k in range(0, len(mods), 3):
        conv, bn, act = mods[k], mods[k + 1], mods[k + 2]
        assert isinstance(conv, nn.Conv2d) and conv.kernel_size == (1, 1) and conv.bias is None
        assert isinstance(bn, nn.BatchNorm2d) and isinstance(act, nn.ReLU)
        w = conv.weight.detach().float().reshape(conv.out_channels, conv.in_channels)
        if not channels:
            channels.append(conv.in_channels)
        channels.append(conv.out_channels)
        gamma = bn.weight.detach().float() if bn.affine else torch.ones_like(bn.running_var)
        beta = bn.bias.detach().float() if bn.affine else torch.zeros_like(bn.running_var)
        scale = gamma / torch.sqrt(bn.running_var.detach().float() + bn.eps)
        shift = beta - bn.running_mean.detach().float() * scale
        parts += [w.t().contiguous().reshape(-1), scale, shift]
    return channels, torch.cat(parts).contiguous()


def tc_supported(channels, nsample, C_in, pool_avg=False):
    """True when the tcgen05 path handles this shape (nsample 32, [3+C_in, 64, 128], max-pool)."""
    ch = (ctypes.c_int * len(channels))(*channels)
    return bool(_lib.lib().mgar_sa_mlp_tc_supported(int(nsample), int(C_in), len(channels) - 1, ctypes.addressof(ch),
                                                    1 if pool_avg else 0))


def tc_group_max_supported(channels, nsample, C_in, pool_avg=False):
    """True when the tcgen05 path can produce the per-group max for this shape WITHOUT the per-centroid features
    (want_out=False): nsample 32 as above, and also 16 / 64 (mgar_sa_mlp_tc_group_max_supported)."""
    ch = (ctypes.c_int * len(channels))(*channels)
    return bool(_lib.lib().mgar_sa_mlp_tc_group_max_supported(int(nsample), int(C_in), len(channels) - 1,
                                                              ctypes.addressof(ch), 1 if pool_avg else 0))


def sa_mlp_forward(channels, params, nsample, xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt, features, idx,
                   empty_mask=None, out=None, pool_avg=False, uniform=None, params_host=None, impl=None,
                   group_max=None, group_size=0, group_zero=None, want_out=True, skip_zero_groups=False,
                   group_live_cnt=None, group_point_cnt=None):
    """-> (M, channels[-1]) fp32.  `idx` as returned by ball_query (or the raw kernel output);
    `uniform=(n, m)` skips the ragged bookkeeping when every cloud has n points and m centroids.

    impl: 'tc' (tcgen05 tensor cores, 3xTF32), 'fp32' (CUDA-core FFMA) or None = tensor cores whenever the
    shape is supported and `params_host` (CPU copy of `params`) is given; env MGAR_SA_MLP overrides.
    `xyz` / `features` may be strided column views of one (N,3+C) array (tc only).
    tc only: `group_max` (ceil(M/group_size), C_out) receives the max over every run of `group_size`
    consecutive centroids (rows with group_zero[g] != 0 stay zero); want_out=False skips the (M, C_out) result;
    skip_zero_groups (with want_out=False) lets the kernel walk only the groups whose flag is 0 - same result, flagged
    groups (empty or zero-padded person slots) cost nothing; group_live_cnt (per group, int32, from
    fastpath.fps_ball_query) additionally restricts the walk to each group's distinct centroids (same maxima);
    group_point_cnt (per group, int32: the cloud holds that many distinct points, slot k = point k mod count, as
    roipoint pooling wraps them) further restricts it to the distinct samples of every ball (same maxima to 1e-6)."""
    need_cuda(new_xyz=new_xyz, idx=idx, params=params, empty_mask=empty_mask, group_max=group_max, group_zero=group_zero,
              group_live_cnt=group_live_cnt, group_point_cnt=group_point_cnt)
    need_dtype(torch.int32, group_live_cnt=group_live_cnt, group_zero=group_zero, group_point_cnt=group_point_cnt)
    need_dtype(torch.float32, xyz=xyz, new_xyz=new_xyz, features=features, params=params)
    need_dtype(torch.int32, idx=idx)
    if not xyz.is_cuda or xyz.dim() != 2 or xyz.stride(1) != 1:
        raise _lib.MgarError("xyz must be a CUDA (N,3) tensor with unit column stride")
    if out is not None and (not out.is_cuda or out.dtype != torch.float32 or out.stride(1) != 1):
        raise _lib.MgarError("out must be a CUDA fp32 tensor with unit column stride")
    M = new_xyz.shape[0]
    B = xyz_batch_cnt.shape[0] if xyz_batch_cnt is not None else xyz.shape[0] // uniform[0]
    C_in = 0 if features is None else features.shape[1]
    if features is not None and (not features.is_cuda or features.dim() != 2 or features.stride(1) != 1):
        raise _lib.MgarError("features must be a CUDA (N,C) tensor with unit column stride")
    if out is None and want_out:
        out = torch.empty((M, channels[-1]), dtype=torch.float32, device=xyz.device)
    if empty_mask is not None:
        empty_mask = empty_mask.to(torch.uint8).contiguous()
    impl = os.environ.get("MGAR_SA_MLP", impl)
    if impl is None:
        impl = 'tc' if (params_host is not None and tc_supported(channels, nsample, C_in, pool_avg)) else 'fp32'
    ch = (ctypes.c_int * len(channels))(*channels)
    un, um = uniform if uniform is not None else (0, 0)
    L = _lib.lib()
    with on_device(xyz):
        if impl == 'tc':
            if params_host is None or params_host.is_cuda or not params_host.is_contiguous():
                raise _lib.MgarError("the tensor-core path needs params_host (contiguous CPU copy of params)")
            skip = out is None and skip_zero_groups and group_max is not None
            ws_bytes = int(L.mgar_sa_mlp_tc_workspace_bytes(M, int(group_size))) if skip else 0
            if uniform is None:                  # ragged batches: room for the prefix arrays (the live walk needs `uniform`)
                ws_bytes = max(ws_bytes, int(L.mgar_stack_workspace_bytes(B)))
            ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=xyz.device) if ws_bytes else None
            _lib.check(L.mgar_sa_mlp_fwd_tc(
                B, M, nsample, C_in, ptr(xyz), xyz.stride(0), ptr(xyz_batch_cnt), ptr(new_xyz), ptr(new_xyz_batch_cnt),
                un, um, ptr(features), 0 if features is None else features.stride(0), ptr(idx), ptr(empty_mask),
                len(channels) - 1, ctypes.addressof(ch), ptr(params), ptr(params_host), ptr(out),
                (-1 if skip_zero_groups else 0) if out is None else out.stride(0), ptr(group_max), int(group_size),
                ptr(group_zero), ptr(group_live_cnt) if skip else 0,
                ptr(group_point_cnt) if (skip and group_live_cnt is not None and group_point_cnt is not None) else 0,
                1 if pool_avg else 0, ptr(ws), ws_bytes, stream_of(xyz)), "sa_mlp_fwd_tc")
        else:
            if not xyz.is_contiguous() or (features is not None and not features.is_contiguous()) or group_max is not None:
                raise _lib.MgarError("the fp32 path needs contiguous xyz / features and has no group_max output")
            ws, nb = _lib.stack_workspace(B, xyz.device) if uniform is None else (None, 0)
            _lib.check(L.mgar_sa_mlp_fwd_fp32(
                B, M, nsample, C_in, ptr(xyz), ptr(xyz_batch_cnt), ptr(new_xyz), ptr(new_xyz_batch_cnt), un, um,
                ptr(features), ptr(idx), ptr(empty_mask), len(channels) - 1, ctypes.addressof(ch), ptr(params),
                ptr(out), out.stride(0), 1 if pool_avg else 0, ptr(ws), nb, stream_of(xyz)), "sa_mlp_fwd")
    return out if want_out else group_max


# ---- training mode: BatchNorm batch statistics + backward, through the C ABI --------------------------------
def train_workspace_bytes(M, nsample, channels, pool_avg=False):
    ch = (ctypes.c_int * len(channels))(*channels)
    return int(_lib.lib().mgar_sa_mlp_train_workspace_bytes_for(int(M), int(nsample), len(channels) - 1,
                                                                ctypes.addressof(ch), 1 if pool_avg else 0))


class SAMlpTrainFunction(torch.autograd.Function):
    """QueryAndGroup + [Conv1x1, BatchNorm (batch statistics), ReLU] x L + pool as ONE autograd node.

    forward(features (N,C_in) or None, weights_flat, gamma_beta_flat, geometry...) -> (out (M,C_last), batch_stats)
    weights_flat = cat_l W_l^T.reshape(-1) ([c_in][c_out]); gamma_beta_flat = cat_l (gamma_l, beta_l);
    batch_stats = cat_l (mean_l, biased var_l), not differentiable.  Gradients: features, weights_flat,
    gamma_beta_flat (the reference's autograd produces the same three; coordinates get none)."""

    @staticmethod
    def forward(ctx, features, weights_flat, gamma_beta_flat, channels, nsample, eps, pool_avg, xyz, xyz_batch_cnt,
                new_xyz, new_xyz_batch_cnt, idx, empty_mask, uniform):
        need_cuda(xyz=xyz, new_xyz=new_xyz, idx=idx, features=features, weights_flat=weights_flat,
                  gamma_beta_flat=gamma_beta_flat, empty_mask=empty_mask)
        need_dtype(torch.float32, xyz=xyz, new_xyz=new_xyz, features=features, weights_flat=weights_flat,
                   gamma_beta_flat=gamma_beta_flat)
        need_dtype(torch.int32, idx=idx)
        M = new_xyz.shape[0]
        C_in = 0 if features is None else features.shape[1]
        B = xyz_batch_cnt.shape[0] if xyz_batch_cnt is not None else xyz.shape[0] // uniform[0]
        un, um = uniform if uniform is not None else (0, 0)
        if empty_mask is not None:
            empty_mask = empty_mask.to(torch.uint8).contiguous()
        ch = (ctypes.c_int * len(channels))(*channels)
        # the training workspace, + the prefix arrays of ragged batches behind it (include/mgar_lidar.h)
        ws_bytes = train_workspace_bytes(M, nsample, channels, pool_avg)
        if uniform is None:
            ws_bytes = ((ws_bytes + 15) & ~15) + int(_lib.lib().mgar_stack_workspace_bytes(B))
        ws = torch.empty((max(ws_bytes, 1),), dtype=torch.uint8, device=xyz.device)
        out = torch.empty((M, channels[-1]), dtype=torch.float32, device=xyz.device)
        stats = torch.empty((2 * sum(channels[1:]),), dtype=torch.float32, device=xyz.device)
        w, gb = weights_flat.detach().contiguous(), gamma_beta_flat.detach().contiguous()
        with on_device(xyz):
            _lib.check(_lib.lib().mgar_sa_mlp_fwd_train(
                B, M, int(nsample), C_in, ptr(xyz), ptr(xyz_batch_cnt), ptr(new_xyz), ptr(new_xyz_batch_cnt), un, um,
                ptr(features), ptr(idx), ptr(empty_mask), len(channels) - 1, ctypes.addressof(ch), ptr(w), ptr(gb),
                float(eps), ptr(out), out.stride(0), 1 if pool_avg else 0, ptr(stats), ptr(ws), ws_bytes,
                stream_of(xyz)), "sa_mlp_fwd_train")
        ctx.save_for_backward(w, gb, idx, xyz_batch_cnt, new_xyz_batch_cnt, empty_mask, ws)
        ctx.geom = (B, M, int(nsample), C_in, un, um, list(channels), bool(pool_avg), ws_bytes,
                    None if features is None else features.shape[0])
        ctx.mark_non_differentiable(stats)
        return out, stats

    @staticmethod
    def backward(ctx, grad_out, _grad_stats):
        w, gb, idx, xyz_batch_cnt, new_xyz_batch_cnt, empty_mask, ws = ctx.saved_tensors
        B, M, nsample, C_in, un, um, channels, pool_avg, ws_bytes, N = ctx.geom
        grad_out = grad_out.contiguous().float()
        ch = (ctypes.c_int * len(channels))(*channels)
        gw, ggb = torch.empty_like(w), torch.empty_like(gb)
        gf = torch.zeros((N, C_in), dtype=torch.float32, device=w.device) if (C_in and ctx.needs_input_grad[0]) else None
        with on_device(w):
            _lib.check(_lib.lib().mgar_sa_mlp_bwd(
                B, M, nsample, C_in, ptr(xyz_batch_cnt), ptr(new_xyz_batch_cnt), un, um, ptr(idx), ptr(empty_mask),
                len(channels) - 1, ctypes.addressof(ch), ptr(w), ptr(gb), ptr(grad_out), grad_out.stride(0),
                1 if pool_avg else 0, ptr(ws), ws_bytes, ptr(gw), ptr(ggb), ptr(gf), stream_of(w)), "sa_mlp_bwd")
        return (gf, gw, ggb) + (None,) * 11


def sa_mlp_train_forward(seq, nsample, xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt, features, idx, empty_mask=None,
                         pool_avg=False, uniform=None):
    """Training-mode forward of `seq` (= [Conv2d 1x1, BatchNorm2d, ReLU] x L, in train()) through the C ABI: returns
    the pooled (M, C_last) features as a differentiable tensor and updates the BatchNorm running statistics exactly
    as nn.BatchNorm2d does (momentum, unbiased running variance, num_batches_tracked)."""
    mods = list(seq)
    convs, bns = mods[0::3], mods[1::3]
    channels = [convs[0].in_channels] + [c.out_channels for c in convs]
    eps = bns[0].eps
    if any(bn.eps != eps or not bn.affine for bn in bns):
        raise _lib.MgarError("training path: BatchNorm layers must be affine and share eps")
    weights_flat = torch.cat([c.weight.reshape(c.out_channels, c.in_channels).t().reshape(-1) for c in convs])
    gamma_beta_flat = torch.cat([t for bn in bns for t in (bn.weight, bn.bias)])
    out, stats = SAMlpTrainFunction.apply(features, weights_flat, gamma_beta_flat, channels, nsample, eps, pool_avg, xyz,
                                          xyz_batch_cnt, new_xyz, new_xyz_batch_cnt, idx, empty_mask, uniform)
    P = new_xyz.shape[0] * nsample
    with torch.no_grad():
        off = 0
        for bn in bns:
            C = bn.num_features
            mean, var = stats[off:off + C], stats[off + C:off + 2 * C]
            off += 2 * C
            if bn.track_running_stats and bn.running_mean is not None:
                bn.num_batches_tracked += 1
                mom = bn.momentum if bn.momentum is not None else 1.0 / float(bn.num_batches_tracked)
                bn.running_mean.mul_(1 - mom).add_(mean, alpha=mom)
                bn.running_var.mul_(1 - mom).add_(var * (P / max(P - 1, 1)), alpha=mom)
    return out
