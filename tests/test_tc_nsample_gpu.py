"""-m gpu: nsample 16 and 64 on the tcgen05 SA-MLP kernel (blocks of 32 positions: two 16-sample centroids or half a
64-sample one, TcArgs::ns) - only the per-group max is produced.  Checked against the exact-fp32 CUDA kernel and the
torch-CPU oracle (1e-5 relative, the north_star's fp32 tolerance), with empty balls, flagged groups, the live-person
walk and the distinct-centroid walk; and end to end through PointPersonEncoder (configs[4]'s sweep points)."""
import numpy as np
import pytest
import torch

from multimodal_gar_b200 import _lib, lidar_branch, synthetic
from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_stack import sa_mlp_fused
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def T(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _close(got, want):
    scale = max(1.0, float(np.abs(want).max()))
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-5 * scale)   # north_star: 1e-5 relative, fp32


@pytest.mark.parametrize("K,Bp,n,m,r", [(16, 9, 256, 64, 0.3), (64, 9, 256, 64, 0.45), (16, 40, 512, 512, 0.1),
                                        (64, 21, 512, 512, 1.0), (64, 5, 128, 6, 0.3), (16, 5, 128, 10, 0.3)])
def test_group_max_nsample_16_64_on_tensor_cores(cuda, K, Bp, n, m, r):
    rng = np.random.default_rng(K * 100 + Bp)
    pooled = rng.uniform(-0.5, 0.5, (Bp * n, 4)).astype(np.float32)
    pooled[:, :3] *= np.array([0.7, 0.7, 1.8], dtype=np.float32)
    xyz, feats = np.ascontiguousarray(pooled[:, :3]), np.ascontiguousarray(pooled[:, 3:4])
    new_xyz = xyz.reshape(Bp, n, 3)[:, np.arange(m) % n].reshape(-1, 3).copy()
    new_xyz[::37] += 9.0                                     # empty balls (first slot -1)
    # distinct-centroid structure: centroids lc[g] .. m-1 of cloud g repeat centroid 0 (what sampling a wrapped cloud gives)
    lc = rng.integers(1, m + 1, Bp).astype(np.int32)
    lc[0], lc[-1] = m, 1
    nz = new_xyz.reshape(Bp, m, 3)
    for g in range(Bp):
        nz[g, lc[g]:] = nz[g, 0]
    cnt, ncnt = np.full(Bp, n, np.int32), np.full(Bp, m, np.int32)
    mlp = O.make_sa_mlp([4, 64, 128], seed=K)
    idx, mask = O.ball_query(r, K, xyz, cnt, new_xyz, ncnt)
    assert mask.any() and not mask.all()
    ch, params = sa_mlp_fused.pack_shared_mlp(mlp)
    assert sa_mlp_fused.tc_group_max_supported(ch, K, 1) and not sa_mlp_fused.tc_supported(ch, K, 1)
    dxyz, dfe, dnew, didx, dmask = T(xyz, cuda), T(feats, cuda), T(new_xyz, cuda), T(idx, cuda), T(mask, cuda)
    ref32 = sa_mlp_fused.sa_mlp_forward(ch, params.to(cuda), K, dxyz, None, dnew, None, dfe, didx, empty_mask=dmask,
                                        impl='fp32', uniform=(n, m))
    want = ref32.view(Bp, m, 128).max(dim=1).values
    zero = np.zeros(Bp, np.int32)
    zero[1] = 1
    want_z = want.clone()
    want_z[1] = 0

    def tc(**kw):
        g = torch.full((Bp, 128), -1.0, device=cuda)
        n0 = _lib.launch_count()
        out = sa_mlp_fused.sa_mlp_forward(ch, params.to(cuda), K, dxyz, None, dnew, None, dfe, didx, empty_mask=dmask,
                                          params_host=params, impl='tc', group_max=g, group_size=m, want_out=False,
                                          uniform=(n, m), **kw)
        torch.cuda.synchronize()
        return out, _lib.launch_count() - n0

    plain, l0 = tc()                                                           # every block
    _close(plain.cpu().numpy(), want.cpu().numpy())
    flagged, l1 = tc(group_zero=T(zero, cuda))
    _close(flagged.cpu().numpy(), want_z.cpu().numpy())
    assert torch.equal(flagged[[0, 2]], plain[[0, 2]]) and l0 == l1 == 1
    blocks_per_group = m * K // 32
    if blocks_per_group % 4 == 0:
        persons, l2 = tc(group_zero=T(zero, cuda), skip_zero_groups=True)     # live-person walk
        tiles, l3 = tc(group_zero=T(zero, cuda), skip_zero_groups=True, group_live_cnt=T(lc, cuda))   # distinct centroids
        assert l2 == 2 and l3 == 2
        assert torch.equal(persons, flagged)
        assert torch.equal(tiles, flagged), "the distinct-centroid walk changed a person's max"
    # raw kernel rows (-1 in slot 0 of an empty ball, no mask): what the fused sampling kernel hands over
    raw = idx.copy()
    raw[mask] = 0
    raw[mask, 0] = -1
    didx_keep, dmask_keep = didx, dmask
    didx, dmask = T(raw, cuda), None
    assert torch.equal(tc()[0], plain)
    didx, dmask = didx_keep, dmask_keep
    # against the torch-CPU oracle of the reference module
    ref = O.stack_sa_module_msg([mlp], [r], [K], xyz, cnt, new_xyz, ncnt, feats).reshape(Bp, m, 128).max(axis=1)
    _close(plain.cpu().numpy(), ref)
    # per-centroid features are not available on this path: the library says so
    with pytest.raises(_lib.MgarError):
        sa_mlp_fused.sa_mlp_forward(ch, params.to(cuda), K, dxyz, None, dnew, None, dfe, didx, empty_mask=dmask,
                                    params_host=params, impl='tc', uniform=(n, m))


@pytest.mark.parametrize("K,r", [(16, 0.1), (64, 1.0), (64, 0.3)])
def test_encoder_nsample_16_64_matches_the_fp32_kernels(cuda, monkeypatch, K, r):
    pts, boxes = synthetic.make_batch(5, 2, N=65536, M=40)
    boxes = np.concatenate([boxes, np.zeros((2, 8, 7), np.float32)], axis=1)      # padded slots: empty persons
    torch.manual_seed(11)
    enc = lidar_branch.PointPersonEncoder(1, 512, 512, r, K, (64, 128)).to(cuda).eval()
    with torch.no_grad():
        for mod in enc.sa.mlps[0]:
            if isinstance(mod, torch.nn.BatchNorm2d):
                mod.running_mean.normal_(0, 0.1)
                mod.running_var.uniform_(0.8, 1.2)
                mod.weight.normal_(1, 0.3)                       # mixed signs do occur: |N(1, 0.3)| rarely, so force a few
                mod.weight[::7] *= -1
                mod.bias.normal_(0, 0.1)
    xyz, feat, bx = T(pts[..., :3], cuda), T(pts[..., 3:4], cuda), T(boxes, cuda)
    stages = []

    def run():
        stages.clear()
        with torch.no_grad():
            out = enc(xyz, feat, bx, mark=lambda name: stages.append(name))
        torch.cuda.synchronize()
        return out

    monkeypatch.delenv("MGAR_TC_NS32_ONLY", raising=False)
    got = run()
    assert "fps_ball_query" in stages and "person_reduce" not in stages, stages   # fused sampling + tensor-core MLP
    monkeypatch.setenv("MGAR_TC_NS32_ONLY", "1")
    want = run()
    assert "person_reduce" in stages, stages                                       # the fp32 composition
    _close(got.cpu().numpy(), want.cpu().numpy())
    assert (got[:, 40:] == 0).all() and got[:, :40].abs().sum() > 0
    enc.skip_repeated_centroids = False
    monkeypatch.delenv("MGAR_TC_NS32_ONLY", raising=False)
    assert torch.equal(run(), got)                                                 # every centroid computed: same bits
