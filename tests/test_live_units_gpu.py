"""-m gpu: the distinct-SAMPLE walk of the tcgen05 SA MLP (TcArgs::live_units): a tile is sixteen octets of sample slots
of one person instead of four whole centroids, and only the octets that start with a distinct sample of the ball are
listed (a ball row repeats its hits once it runs out of distinct ones: copies of the wrapped cloud, then the first hit
as padding, ball_query_gpu.cu:49-60 / roipoint_pool3d_kernel.cu:86-97).  The per-person max must equal the full walk's
(every centroid, every sample) to the rounding of the operand split (the power-of-two scale of a 32-position block is
shared by other rows than in the full walk: not the same bits, <= 2e-6 relative), the reference's nn.Sequential on
torch-CPU at 1e-5, and the distinct-centroid (tile) walk must keep its bits."""
import numpy as np
import pytest
import torch

from multimodal_gar_b200 import _lib, fastpath, lidar_branch, synthetic
from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_stack import sa_mlp_fused
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def T(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _close(got, want, tol):
    scale = max(1.0, float(np.abs(want).max()))
    np.testing.assert_allclose(got, want, rtol=tol, atol=tol * scale)


def _enc(dev, seed=7):
    torch.manual_seed(seed)
    enc = lidar_branch.PointPersonEncoder(1).to(dev).eval()
    with torch.no_grad():
        for mod in enc.sa.mlps[0]:
            if isinstance(mod, torch.nn.BatchNorm2d):
                mod.running_mean.normal_(0, 0.1)
                mod.running_var.uniform_(0.8, 1.2)
                mod.weight.normal_(0.3, 1.0)          # mixed-sign BatchNorm scales (the epilogue folds the sign into W2)
                mod.bias.normal_(0, 0.1)
    return enc


def _person(enc, pts, boxes, dev, units, tiles=True):
    enc.skip_repeated_samples, enc.skip_repeated_centroids = units, tiles
    xyz, feat = T(pts[..., :3], dev), T(pts[..., 3:4], dev)
    n0 = _lib.launch_count()
    with torch.no_grad():
        out = enc(xyz, feat, T(boxes, dev))
    torch.cuda.synchronize()
    return out, _lib.launch_count() - n0


@pytest.mark.parametrize("case,density", [("mixed", 0.6), ("dense", 4.0), ("sparse", 0.05), ("padded", 1.0), ("all_empty", 1.0)])
def test_unit_walk_equals_the_full_walk(cuda, case, density):
    B, M = 3, 40
    pts, boxes = synthetic.make_batch(77, B, N=32768, M=M, density_scale=density)
    if case == "padded":
        boxes = np.concatenate([boxes, np.zeros((B, 24, 7), np.float32)], axis=1)
    if case == "all_empty":
        boxes[..., :2] += 500.0
    enc = _enc(cuda)
    units, n_u = _person(enc, pts, boxes, cuda, True)
    tiles, n_t = _person(enc, pts, boxes, cuda, False)
    full, n_f = _person(enc, pts, boxes, cuda, False, tiles=False)
    assert n_u == n_t == n_f >= 6, (n_u, n_t, n_f)            # one list-building launch each (units / tiles / live persons)
    assert torch.equal(tiles, full), "the distinct-centroid walk changed bits"
    _close(units.cpu().numpy(), full.cpu().numpy(), 2e-6)
    assert torch.equal(units == 0, full == 0)                  # ReLU zeros and empty persons are exact zeros in both
    if case == "all_empty":
        assert (units == 0).all()
    if case == "padded":
        assert (units[:, 40:] == 0).all() and units[:, :40].abs().sum() > 0


def test_unit_walk_against_torch_cpu_on_sampled_persons(cuda):
    B, M, S, P, K, r = 2, 40, 512, 512, 32, 0.3
    pts, boxes = synthetic.make_batch(5, B, N=65536, M=M)
    enc = _enc(cuda, seed=11)
    units, _ = _person(enc, pts, boxes, cuda, True)
    with torch.no_grad():
        st = enc(T(pts[..., :3], cuda), T(pts[..., 3:4], cuda), T(boxes, cuda), return_stages=True)
    pooled = st["pooled"].view(B * M, S, 4).cpu().numpy()
    fps = st["fps_idx"].cpu().numpy()
    empty = st["empty"].view(-1).cpu().numpy()
    import copy
    mlp = copy.deepcopy(enc.sa.mlps[0]).cpu()
    got = units.view(-1, 128).cpu().numpy()
    cnt = np.array([int(len(np.unique(pooled[g], axis=0))) for g in range(B * M)])
    live = np.flatnonzero(empty == 0)
    # the smallest, a mid-sized and the largest clouds, plus a few random ones
    order = live[np.argsort(cnt[live])]
    pick = sorted(set(order[:2].tolist()) | set(order[-2:].tolist()) | {int(order[len(order) // 2])} |
                  set(np.random.default_rng(3).choice(live, 4, replace=False).tolist()))
    for g in pick:
        cxyz = np.ascontiguousarray(pooled[g, :, :3])
        want = O.stack_sa_module_msg([mlp], [r], [K], cxyz, np.array([S], np.int32), cxyz[fps[g]], np.array([P], np.int32),
                                     np.ascontiguousarray(pooled[g, :, 3:])).max(axis=0)
        _close(got[g], want, 1e-5)
    assert (got[empty != 0] == 0).all()


@pytest.mark.parametrize("n,m,periods", [(128, 128, (128, 1, 7, 32, 33, 100)), (512, 512, (512, 511, 8, 9, 64, 300))])
def test_unit_walk_through_the_c_abi_with_given_periods(cuda, n, m, periods):
    """Clouds built here (not pooled): cloud g repeats its first periods[g] points; centroids = the sampling of it.  The
    count passed as group_point_cnt is the contract the walk relies on; with NULL the tile walk runs."""
    K, r = 32, 0.25
    rng = np.random.default_rng(n)
    Bp = len(periods)
    base = rng.uniform(-0.5, 0.5, (Bp, n, 4)).astype(np.float32)
    base[..., :3] *= np.array([0.6, 0.6, 1.2], dtype=np.float32)
    for g, p in enumerate(periods):
        base[g] = base[g, np.arange(n) % p]
    pooled = T(base, cuda)
    fused = fastpath.fps_ball_query(pooled[:, :, 0:3], m, r, K, all_rows=True)
    assert fused is not None
    fps_idx, new_xyz, ball_idx, live_cnt = fused
    assert live_cnt.cpu().tolist() == [min(p, m) for p in periods]
    mlp = O.make_sa_mlp([4, 64, 128], seed=3)
    ch, params = sa_mlp_fused.pack_shared_mlp(mlp)
    flat = pooled.view(Bp * n, 4)
    zero = torch.zeros((Bp,), dtype=torch.int32, device=cuda)

    def run(**kw):
        g = torch.full((Bp, 128), -1.0, device=cuda)
        return sa_mlp_fused.sa_mlp_forward(ch, params.to(cuda), K, flat[:, 0:3], None, new_xyz.view(-1, 3), None, flat[:, 3:],
                                           ball_idx, uniform=(n, m), params_host=params, impl='tc', group_max=g, group_size=m,
                                           group_zero=zero, want_out=False, skip_zero_groups=True, **kw)

    full = run()
    tiles = run(group_live_cnt=live_cnt)
    units = run(group_live_cnt=live_cnt, group_point_cnt=T(np.array(periods, np.int32), cuda))
    torch.cuda.synchronize()
    assert torch.equal(tiles, full)
    _close(units.cpu().numpy(), full.cpu().numpy(), 2e-6)
    # the reference module on torch-CPU
    cnt, ncnt = np.full(Bp, n, np.int32), np.full(Bp, m, np.int32)
    ref = O.stack_sa_module_msg([mlp], [r], [K], np.ascontiguousarray(base[..., :3]).reshape(-1, 3), cnt,
                                new_xyz.view(-1, 3).cpu().numpy(), ncnt,
                                np.ascontiguousarray(base[..., 3:]).reshape(-1, 1)).reshape(Bp, m, 128).max(axis=1)
    _close(units.cpu().numpy(), ref, 1e-5)
