"""Where the fused DAFM step spends its time (64 frames x 100 persons, D = 512): projection GEMMs (side by side vs
separate), the attention kernel alone, the FFN GEMMs, the add + LayerNorm kernel, and the torch attention core."""
import ctypes
import json
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from multimodal_gar_b200 import _lib, dafm  # noqa: E402
from scripts.time_dafm import timed  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    B, n, D = 64, 100, 512
    mod = dafm.FusionAttentionMat(D, D).to(dev).eval()
    R, L = torch.randn(B, n, D, device=dev), torch.randn(B, n, D, device=dev)
    E = torch.softmax(-torch.rand(B, n, n, device=dev), dim=-1)
    valid = torch.ones(B, n, dtype=torch.bool, device=dev)
    res = {}
    with torch.no_grad():
        W3 = torch.cat([mod.WQ_r, mod.WK_l, mod.WV_l], dim=1)
        res["cat_ms"] = timed(lambda: torch.cat([mod.WQ_r, mod.WK_l, mod.WV_l], dim=1))
        res["gemm_1536_ms"] = timed(lambda: L.view(B * n, D) @ W3)
        res["gemm_3x512_ms"] = timed(lambda: (L.view(B * n, D) @ mod.WQ_r, L.view(B * n, D) @ mod.WK_l, L.view(B * n, D) @ mod.WV_l))
        PL, PR = L.view(B * n, D) @ W3, R.view(B * n, D) @ W3
        X = torch.empty(2, B, n, D, device=dev)
        lib = _lib.lib()
        st = torch.cuda.current_stream().cuda_stream
        arr = lambda a, b: (ctypes.c_void_p * 2)(a, b)
        pl, pr, f4 = PL.data_ptr(), PR.data_ptr(), 4 * D

        def att():
            _lib.check(lib.mgar_dafm_attention(B, n, D, 2, 3 * D, arr(pl, pr + 2 * f4), arr(pr, pl + f4), arr(pr + f4, pl + 2 * f4),
                                               arr(R.data_ptr(), L.data_ptr()), E.data_ptr(), valid.data_ptr(),
                                               arr(mod.LN_r_1.weight.data_ptr(), mod.LN_l_1.weight.data_ptr()),
                                               arr(mod.LN_r_1.bias.data_ptr(), mod.LN_l_1.bias.data_ptr()), 1e-5,
                                               arr(X[0].data_ptr(), X[1].data_ptr()), st), "att")
        res["attention_kernel_2streams_ms"] = timed(att)
        x = X[0].view(B * n, D)
        res["ffn_ms"] = timed(lambda: mod.FFN_r(x))
        y = mod.FFN_r(x)
        o = torch.empty_like(x)
        res["add_ln_ms"] = timed(lambda: lib.mgar_dafm_add_layernorm(B * n, D, x.data_ptr(), y.data_ptr(), mod.LN_r_2.weight.data_ptr(),
                                                                     mod.LN_r_2.bias.data_ptr(), valid.data_ptr(), 1e-5, o.data_ptr(), st))
        Q, K, V = PL[:, :D].view(B, n, D), PR[:, :D].view(B, n, D), PR[:, D:2 * D].view(B, n, D)

        def core():
            a = torch.softmax((Q @ K.transpose(-1, -2)) * E / D ** 0.5, dim=-1)
            return mod.LN_r_1(a @ V + R)
        res["torch_attention_core_1stream_ms"] = timed(core)
        res["torch_add_ln_ms"] = timed(lambda: mod.LN_r_2(x + y))
    print(json.dumps(res))
    json.dump(res, open("gpurun_out/dafm_diag.json", "w"))


if __name__ == "__main__":
    os.makedirs("gpurun_out", exist_ok=True)
    main()
