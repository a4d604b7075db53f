"""fp32-faithful linear layers on tcgen05 (libmgar_lidar.so: csrc/gemm_tc.cu, 3xTF32) for the dense products around the
LiDAR hot path that the reference leaves to torch / cuBLAS fp32: the projections and FFN linears of
`FusionAttention_mat` (/root/reference/model/gat_model.py:466-511) and the 128 -> 512 LiDAR embedding (:883-894).

Inference only (no autograd node): callers keep torch's own `F.linear` whenever gradients are required, the tensors are
not CUDA fp32, or the shape is outside the kernel's (out_features % 128 == 0, in_features % 32 == 0)."""
import weakref

import torch

from . import _lib
from ._torch_glue import on_device, ptr, stream_of


def usable(x, out_features, in_features):
    """True when `x` (.., in_features) can go through the tensor-core kernel right now."""
    if not (x.is_cuda and x.dtype == torch.float32) or torch.is_grad_enabled() and x.requires_grad:
        return False
    rows = x.numel() // max(in_features, 1)
    return rows > 0 and bool(_lib.lib().mgar_gemm_tf32x3_supported(int(rows), int(out_features), int(in_features)))


_PACKED = {}          # (data_ptr, version, shape, strides, device) -> packed weight tiles (mgar_gemm_pack_bt), small LRU


def packed_weight(w):
    """The weight (N, K) split into tf32 hi / lo and laid out tile by tile as the kernel reads it (csrc/gemm_tc.cu);
    prepared once per weight VALUE: the key holds the storage pointer and the in-place version counter."""
    key = (w.data_ptr(), w._version, tuple(w.shape), tuple(w.stride()), str(w.device))
    hit = _PACKED.get(key)
    if hit is not None and hit[0]() is not w:          # another tensor now lives at that address: not the same weight
        hit = None
    if hit is not None:
        return hit[1]
    N, K = w.shape
    src = w.detach()
    if src.stride(1) != 1 or src.stride(0) % 4 or src.data_ptr() % 16:
        src = src.contiguous()
    L = _lib.lib()
    hit = torch.empty((int(L.mgar_gemm_packed_bytes(N, K)) // 4,), dtype=torch.float32, device=w.device)
    with on_device(w):
        _lib.check(L.mgar_gemm_pack_bt(N, K, ptr(src), src.stride(0), ptr(hit), stream_of(w)), "gemm_pack_bt")
    if len(_PACKED) >= 64:
        _PACKED.pop(next(iter(_PACKED)))
    _PACKED[key] = (weakref.ref(w), hit)
    return hit


def linear(x, weight, bias=None, relu=False, out=None, packed=True):
    """act(x @ weight.T + bias) with `weight` (out_features, in_features) as nn.Linear stores it.  x (.., in) fp32 CUDA;
    returns (.., out).  Raises MgarError when the shape is unsupported (ask `usable` first).  packed=True (default): the
    weight goes through `packed_weight` (prepared once, bulk-copied by the kernel); False: the kernel splits it per call."""
    K, N = weight.shape[1], weight.shape[0]
    x2 = x.reshape(-1, K)
    if x2.stride(1) != 1 or x2.stride(0) % 4 or x2.data_ptr() % 16:
        x2 = x2.contiguous()
    w = weight if (weight.stride(1) == 1 and weight.stride(0) % 4 == 0 and weight.data_ptr() % 16 == 0) else weight.contiguous()
    M = x2.shape[0]
    y = out if out is not None else torch.empty((M, N), dtype=torch.float32, device=x.device)
    if y.stride(1) != 1 or y.stride(0) % 4 or y.data_ptr() % 16:
        raise _lib.MgarError("linear: output must be a 16-byte aligned row-major fp32 tensor")
    b = None if bias is None else bias.detach().contiguous()
    with on_device(x):
        if packed:
            pw = packed_weight(weight)
            _lib.check(_lib.lib().mgar_gemm_tf32x3_packed(M, N, K, ptr(x2), x2.stride(0), ptr(pw), ptr(b), 1 if relu else 0,
                                                          ptr(y), y.stride(0), stream_of(x)), "gemm_tf32x3_packed")
        else:
            _lib.check(_lib.lib().mgar_gemm_tf32x3(M, N, K, ptr(x2), x2.stride(0), ptr(w.detach()), w.stride(0), ptr(b),
                                                   1 if relu else 0, ptr(y), y.stride(0), stream_of(x)), "gemm_tf32x3")
    return y if out is not None else y.view(*x.shape[:-1], N)


def linear_grouped(items, relu=False):
    """[act(x_g @ W_g.T + b_g)] for up to four (x, weight, bias) triples of ONE shape in one launch
    (mgar_gemm_tf32x3_packed_grouped): the tiles of all products fill the SMs together.  Falls back to `linear` per item
    when the shapes differ."""
    import ctypes
    if len(items) == 1 or len(items) > 4 or len({(tuple(x.shape), tuple(w.shape)) for x, w, _ in items}) != 1:
        return [linear(x, w, b, relu=relu) for x, w, b in items]
    K, N = items[0][1].shape[1], items[0][1].shape[0]
    xs, ys, pws, bs = [], [], [], []
    for x, w, b in items:
        x2 = x.reshape(-1, K)
        if x2.stride(1) != 1 or x2.stride(0) % 4 or x2.data_ptr() % 16:
            x2 = x2.contiguous()
        xs.append(x2)
        ys.append(torch.empty((x2.shape[0], N), dtype=torch.float32, device=x.device))
        pws.append(packed_weight(w))
        bs.append(None if b is None else b.detach().contiguous())
    M, G = xs[0].shape[0], len(items)
    vp, ip = ctypes.c_void_p * G, ctypes.c_int * G
    a_ = vp(*[ptr(t) for t in xs]); p_ = vp(*[ptr(t) for t in pws]); b_ = vp(*[ptr(t) for t in bs]); c_ = vp(*[ptr(t) for t in ys])
    lda = ip(*[t.stride(0) for t in xs]); ldc = ip(*[t.stride(0) for t in ys])
    with on_device(xs[0]):
        _lib.check(_lib.lib().mgar_gemm_tf32x3_packed_grouped(G, M, N, K, ctypes.addressof(a_), ctypes.addressof(lda),
                                                              ctypes.addressof(p_), ctypes.addressof(b_), 1 if relu else 0,
                                                              ctypes.addressof(c_), ctypes.addressof(ldc), stream_of(xs[0])),
                   "gemm_tf32x3_packed_grouped")
    return [y.view(*x.shape[:-1], N) for y, (x, _, _) in zip(ys, items)]


class CachedTranspose:
    """(K, N) parameters used as `x @ W` -> the (N, K) K-major copy the kernel wants, re-made when a parameter changes
    (storage pointer + in-place version, as the folded-BatchNorm cache of the SA module)."""

    def __init__(self):
        self.key, self.value = None, None

    def get(self, params):
        key = tuple((p.data_ptr(), p._version, str(p.device)) for p in params)
        if key != self.key:
            with torch.no_grad():
                self.value = torch.cat([p.detach() for p in params], dim=1).t().contiguous()
            self.key = key
        return self.value
