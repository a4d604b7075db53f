"""-m gpu: the CUDA path (libmgar_lidar.so through the pcdet.ops mirrors) against the CPU oracle on
the same seeded inputs.  Bit-exact for masks / indices / gathered values; 1e-5 relative (of the
feature scale) for the fp32 shared-MLP features and the canonical rotation."""
import numpy as np
import pytest
import torch

import cases
from multimodal_gar_b200 import _lib, synthetic
from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_stack import pointnet2_modules, pointnet2_utils, sa_mlp_fused
from multimodal_gar_b200.pcdet.ops.roiaware_pool3d import roiaware_pool3d_utils
from multimodal_gar_b200.pcdet.ops.roipoint_pool3d import roipoint_pool3d_cuda, roipoint_pool3d_utils
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def T(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _assert_in_box_equal(got, want, xyz, boxes):
    """glibc (oracle) vs libdevice (GPU) cosf/sinf may differ by an ulp: only pairs whose decision is
    sensitive to that may differ (the reference-GPU tests compare every pair bit-exactly)."""
    bad = got != want
    if bad.any():
        amb = np.stack([O.in_box_ambiguous(xyz[b], boxes[b]).any(axis=0) for b in range(xyz.shape[0])])
        assert not (bad & ~amb).any(), "%d unambiguous in-box decisions differ" % int((bad & ~amb).sum())


@pytest.mark.parametrize("N,M", [(4096, 12), (4099, 7), (1, 1), (65536, 40), (513, 1100)])
def test_points_in_boxes_gpu(cuda, N, M):
    pts, boxes = synthetic.make_batch(5, 2, N=N, M=M, density_scale=0.3)
    xyz = np.ascontiguousarray(pts[..., :3])
    got = roiaware_pool3d_utils.points_in_boxes_gpu(T(xyz, cuda), T(boxes, cuda)).cpu().numpy()
    assert got.dtype == np.int32 and got.shape == (2, N)
    _assert_in_box_equal(got, O.points_in_boxes_gpu(xyz, boxes), xyz, boxes)


def test_points_in_boxes_gpu_faces_and_padding(cuda):
    xyz, _, boxes = cases.frame_case()
    got = roiaware_pool3d_utils.points_in_boxes_gpu(T(xyz, cuda), T(boxes, cuda)).cpu().numpy()
    want = O.points_in_boxes_gpu(xyz, boxes)
    _assert_in_box_equal(got, want, xyz, boxes)
    M = boxes.shape[1]
    assert got[1, 0] == M - 2 and got[1, 1] == M - 2 and got[1, 2] != M - 2


def test_points_in_boxes_gpu_no_boxes(cuda):
    out = roiaware_pool3d_utils.points_in_boxes_gpu(torch.zeros(1, 100, 3, device=cuda),
                                                    torch.zeros(1, 0, 7, device=cuda))
    assert (out == -1).all()


@pytest.mark.parametrize("as_numpy", [True, False])
def test_points_in_boxes_cpu_api_on_gpu(cuda, as_numpy):
    """The host-buffer entry: margin 1e-2 semantics, bit-identical to the reference CPU function."""
    pts, boxes = synthetic.make_frame(1994, N=16384, M=40)
    xyz = np.ascontiguousarray(pts[:, :3])
    want = O.points_in_boxes_cpu(xyz, boxes)
    if as_numpy:
        got = roiaware_pool3d_utils.points_in_boxes_cpu(xyz, boxes)
        assert isinstance(got, np.ndarray)
    else:
        got = roiaware_pool3d_utils.points_in_boxes_cpu(torch.from_numpy(xyz), torch.from_numpy(boxes)).numpy()
    np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("extra,S", [((0.0, 0.0, 0.0), 128), ((0.2, 0.2, 0.2), 64), ((0.0, 0.0, 0.0), 512),
                                     ((1.0, 1.0, 1.0), 2048)])
def test_roipoint_pool3d(cuda, extra, S):
    xyz, feat, boxes = cases.frame_case()
    layer = roipoint_pool3d_utils.RoIPointPool3d(num_sampled_points=S, pool_extra_width=extra)
    pooled, empty = layer(T(xyz, cuda), T(feat, cuda), T(boxes, cuda))
    want_p, want_e, want_i = O.roipoint_pool3d(xyz, feat, boxes, extra, S, return_idx=True)
    assert pooled.shape == want_p.shape and empty.dtype == torch.int32
    if not np.array_equal(empty.cpu().numpy(), want_e) or not np.array_equal(pooled.cpu().numpy(), want_p):
        # only ulp-ambiguous pairs may change membership
        big = O.enlarge_box3d(boxes, extra)
        amb = sum(int(O.in_box_ambiguous(xyz[b], big[b]).sum()) for b in range(xyz.shape[0]))
        assert amb > 0, "pooled output differs from the oracle with no ambiguous pair"
    assert (want_e == 0).sum() > 0 and (want_e.sum() > 0 or extra[0] >= 1.0)      # both branches exercised


def test_roipoint_pool3d_multi_channel_and_idx(cuda):
    xyz, _, boxes = cases.frame_case()
    rng = np.random.default_rng(3)
    feat = rng.random((xyz.shape[0], xyz.shape[1], 5)).astype(np.float32)
    B, M, S = xyz.shape[0], boxes.shape[1], 96
    pooled = torch.zeros((B, M, S, 8), device=cuda)
    empty = torch.zeros((B, M), dtype=torch.int32, device=cuda)
    pidx = torch.zeros((B, M, S), dtype=torch.int32, device=cuda)
    roipoint_pool3d_cuda.forward(T(xyz, cuda), T(boxes, cuda), T(feat, cuda), pooled, empty, pts_idx=pidx)
    want_p, want_e, want_i = O.roipoint_pool3d(xyz, feat, boxes, (0, 0, 0), S, return_idx=True)
    np.testing.assert_array_equal(empty.cpu().numpy(), want_e)
    np.testing.assert_array_equal(pooled.cpu().numpy(), want_p)
    live = want_e == 0
    np.testing.assert_array_equal(pidx.cpu().numpy()[live], want_i[live])


def test_roipoint_pool3d_fused_canonical(cuda):
    xyz, feat, boxes = cases.frame_case()
    B, M, S = xyz.shape[0], boxes.shape[1], 128
    pooled = torch.zeros((B, M, S, 4), device=cuda)
    empty = torch.zeros((B, M), dtype=torch.int32, device=cuda)
    roipoint_pool3d_cuda.forward(T(xyz, cuda), T(boxes, cuda), T(feat, cuda), pooled, empty, canonical=True)
    want_p, want_e = O.roipoint_pool3d(xyz, feat, boxes, (0, 0, 0), S)
    want_c = O.canonical_transform(want_p, boxes, want_e)
    np.testing.assert_array_equal(empty.cpu().numpy(), want_e)
    got = pooled.cpu().numpy().reshape(B * M, S, 4)
    np.testing.assert_allclose(got[..., :3], want_c[..., :3], rtol=0, atol=1e-5 * 2.0)   # 1e-5 of the box scale
    np.testing.assert_array_equal(got[..., 3], want_c[..., 3])


@pytest.mark.parametrize("N,M,S,extra", [(4096, 12, 128, 0.0), (4099, 12, 64, 0.2), (65536, 40, 512, 0.0), (1000, 3, 16, 1.0),
                                         (33, 2, 8, 0.0)])
def test_mask_pooling_equals_single_kernel_pooling_and_points_in_boxes(cuda, N, M, S, extra):
    """Two-pass pooling (box_masks + roipoint_pool_from_masks) against the one-kernel path, the first-hit kernel and
    the oracle: bit masks, box_idx, pooled rows (canonical and raw), empty flags - all bit-exact."""
    from multimodal_gar_b200 import fastpath
    from multimodal_gar_b200.pcdet.utils import box_utils
    pts, boxes = synthetic.make_batch(N + M, 2, N=N, M=M, density_scale=0.5 if N < 60000 else 1.0, empty_frac=0.15)
    xyz, feat = T(pts[..., :3], cuda), T(pts[..., 3:4], cuda)
    bx = T(boxes, cuda)
    big = box_utils.enlarge_box3d(bx.reshape(-1, 7), (extra, extra, extra)).view(2, M, 7).contiguous()
    box_idx, mask = fastpath.box_masks(xyz, bx, None if extra == 0.0 else big)
    assert torch.equal(box_idx, roiaware_pool3d_utils.points_in_boxes_gpu(xyz, bx))
    # masks against the (M,N) in-box matrix of the GPU semantics
    x_np = np.ascontiguousarray(pts[..., :3])
    big_np = big.cpu().numpy()
    bits = np.unpackbits(mask.cpu().numpy().view(np.uint8).reshape(2, M, -1), axis=-1, bitorder="little")[..., :N]
    for b in range(2):
        want = O.in_box_mask_gpu_sem(x_np[b], big_np[b])
        amb = O.in_box_ambiguous(x_np[b], big_np[b])
        assert ((bits[b] != want) & ~amb.astype(bool)).sum() == 0
    for canonical in (True, False):
        pooled, empty = fastpath.roipoint_pool_from_masks(xyz, feat, big, mask, S, canonical=canonical)
        p2 = torch.zeros((2, M, S, 4), device=cuda)
        e2 = torch.zeros((2, M), dtype=torch.int32, device=cuda)
        roipoint_pool3d_cuda.forward(xyz, big, feat, p2, e2, canonical=canonical)
        assert torch.equal(empty, e2)
        assert torch.equal(pooled, p2)


@pytest.mark.parametrize("N", [65536, 3072, 1024])
def test_box_masks_bulk_staged_tiles_equal_the_per_thread_loads(cuda, monkeypatch, N):
    """MGAR_BOX_MASK_BULK=1: the point tiles travel by cp.async.bulk + mbarrier (opt-in; measured no faster) - same
    first-hit indices and bit masks, incl. an odd number of tiles per frame (the CTA's second tile is absent)."""
    from multimodal_gar_b200 import fastpath
    pts, boxes = synthetic.make_batch(N, 3, N=N, M=40, density_scale=0.5 if N < 60000 else 1.0)
    xyz, bx = T(pts[..., :3], cuda), T(boxes, cuda)
    monkeypatch.delenv("MGAR_BOX_MASK_BULK", raising=False)
    idx0, mask0 = fastpath.box_masks(xyz, bx, None)
    monkeypatch.setenv("MGAR_BOX_MASK_BULK", "1")
    idx1, mask1 = fastpath.box_masks(xyz, bx, None)
    assert torch.equal(idx0, idx1) and torch.equal(mask0, mask1)
    assert int((idx0 >= 0).sum()) > 0


def test_mask_pooling_box_grid_adversarial(cuda):
    """The box-grid pass must never lose a hit: boxes stacked in one cell (overflow -> exhaustive loop), boxes and points
    beyond the grid's +-32 m (border cells), boxes straddling cell borders, a huge box, a zero-padded box."""
    from multimodal_gar_b200 import fastpath
    rng = np.random.default_rng(3)
    N = 20000
    frames = []
    # frame 0: twelve boxes on the same spot (cell overflow)
    b0 = np.tile(np.array([[3.0, 4.0, 0.0, 0.8, 0.7, 1.8, 0.3]], np.float32), (12, 1))
    b0[:, :2] += rng.normal(0, 0.05, (12, 2)).astype(np.float32)
    # frame 1: boxes far outside the grid, on cell borders, one huge box, one zero box
    b1 = np.array([[40.0, -45.0, 0.0, 0.9, 0.6, 1.7, 1.0], [-33.0, 31.9, 0.0, 0.9, 0.9, 1.7, -2.0], [5.0, 5.0, 0.0, 0.6, 0.6, 1.6, 0.0],
                   [0.0, 0.0, 0.0, 30.0, 20.0, 3.0, 0.4], [0, 0, 0, 0, 0, 0, 0], [-31.99, -32.0, 0.2, 1.0, 1.0, 1.9, 0.7],
                   [12.0, -7.0, 0.1, 0.5, 0.9, 1.5, 2.5], [12.3, -7.2, 0.1, 0.5, 0.9, 1.5, 2.6], [1e3, 1e3, 0.0, 0.5, 0.5, 1.5, 0.0],
                   [2.0, 2.0, 0.0, 0.5, 0.5, 1.5, 0.0], [2.0, 3.0, 0.0, 0.5, 0.5, 1.5, 0.0], [3.0, 2.0, 0.0, 0.5, 0.5, 1.5, 0.0]], np.float32)
    for boxes in (b0, b1):
        pts = rng.uniform(-50, 50, (N, 3)).astype(np.float32)
        pts[:, 2] = rng.uniform(-1, 1, N)
        for k, bx in enumerate(boxes):                          # seed points inside / around every box
            n = 150
            loc = rng.uniform(-0.6, 0.6, (n, 3)).astype(np.float32) * np.maximum(bx[3:6], 0.1)
            c, s_ = np.cos(bx[6]), np.sin(bx[6])
            pts[k * n:(k + 1) * n, 0] = bx[0] + loc[:, 0] * c - loc[:, 1] * s_
            pts[k * n:(k + 1) * n, 1] = bx[1] + loc[:, 0] * s_ + loc[:, 1] * c
            pts[k * n:(k + 1) * n, 2] = bx[2] + loc[:, 2]
        frames.append((pts, boxes))
    xyz = T(np.stack([f[0] for f in frames]), cuda)
    bx = T(np.stack([f[1] for f in frames]), cuda)
    feat = T(rng.random((2, N, 1)).astype(np.float32), cuda)
    box_idx, mask = fastpath.box_masks(xyz, bx)
    assert torch.equal(box_idx, roiaware_pool3d_utils.points_in_boxes_gpu(xyz, bx))
    assert (box_idx >= 0).sum() > 1000
    pooled, empty = fastpath.roipoint_pool_from_masks(xyz, feat, bx, mask, 64, canonical=False)
    p2 = torch.zeros((2, 12, 64, 4), device=cuda)
    e2 = torch.zeros((2, 12), dtype=torch.int32, device=cuda)
    roipoint_pool3d_cuda.forward(xyz, bx, feat, p2, e2)
    assert torch.equal(empty, e2) and torch.equal(pooled, p2)
    import os
    os.environ["MGAR_NO_BOX_GRID"] = "1"                        # the exhaustive pass gives the same masks
    try:
        _, mask2 = fastpath.box_masks(xyz, bx)
    finally:
        del os.environ["MGAR_NO_BOX_GRID"]
    assert torch.equal(mask, mask2)


@pytest.mark.parametrize("n,m", [(1, 1), (31, 31), (128, 64), (200, 64), (300, 300), (512, 512), (777, 100),
                                 (1024, 128), (1500, 96), (2500, 40), (5000, 64), (9000, 33), (20000, 16)])
def test_fps_batch_with_ties(cuda, n, m):
    c = cases.cloud_with_ties(n, 3, n)
    got = pointnet2_utils.farthest_point_sample(T(c, cuda), m).cpu().numpy()
    np.testing.assert_array_equal(got, O.farthest_point_sample(c, m))


@pytest.mark.parametrize("b,n,m", [(600, 512, 512), (600, 300, 100), (640, 77, 77), (600, 33, 8)])
def test_fps_many_small_clouds(cuda, b, n, m):
    """Many clouds route to the warp-per-cloud kernel; wrapped clouds exercise its early exit (max distance 0)."""
    c = cases.cloud_with_ties(n + 1, b, n)
    c[::3] = np.random.default_rng(b).uniform(-1, 1, c[::3].shape).astype(np.float32)   # every third cloud: no duplicates
    from multimodal_gar_b200 import fastpath
    got = pointnet2_utils.farthest_point_sample(T(c, cuda), m).cpu().numpy()
    pick = list(range(0, b, 25))
    want = O.farthest_point_sample(np.ascontiguousarray(c[pick]), m)
    np.testing.assert_array_equal(got[pick], want)
    # strided in-place variant with the fused gather (what the fused branch calls)
    c4 = np.concatenate([c, np.zeros((b, n, 1), np.float32)], axis=2)
    t4 = T(c4, cuda)
    idx, new_xyz = fastpath.fps_gather(t4[:, :, 0:3], m)
    np.testing.assert_array_equal(idx.cpu().numpy(), got)
    np.testing.assert_array_equal(new_xyz.cpu().numpy(), np.take_along_axis(c, got[..., None].astype(np.int64), axis=1))


def test_fps_batch_temp_contract(cuda):
    from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_stack import pointnet2_stack_cuda as ext
    c = cases.cloud_with_ties(9, 4, 600)
    temp = torch.full((4, 600), 1e10, device=cuda)
    idx = torch.zeros((4, 50), dtype=torch.int32, device=cuda)
    ext.farthest_point_sampling_wrapper(4, 600, 50, T(c, cuda), temp, idx)
    want_i, want_t = O.farthest_point_sample(c, 50, return_temp=True)
    np.testing.assert_array_equal(idx.cpu().numpy(), want_i)
    np.testing.assert_array_equal(temp.cpu().numpy(), want_t)


def test_fps_stack_ragged(cuda):
    xyz, cnt, _, _, _ = cases.ragged_case()
    pos = cnt > 0
    npoint = [32, 64, 16, 40]
    got = pointnet2_utils.stack_farthest_point_sample(T(xyz, cuda), T(cnt[pos], cuda), npoint).cpu().numpy()
    np.testing.assert_array_equal(got, O.stack_farthest_point_sample(xyz, cnt[pos], npoint))
    got = pointnet2_utils.stack_farthest_point_sample(T(xyz, cuda), T(cnt[pos], cuda), 20).cpu().numpy()
    np.testing.assert_array_equal(got, O.stack_farthest_point_sample(xyz, cnt[pos], 20))


def test_fps_stack_large_cloud(cuda):
    rng = np.random.default_rng(11)
    cnt = np.array([12000, 3000], dtype=np.int32)
    xyz = rng.uniform(-5, 5, (int(cnt.sum()), 3)).astype(np.float32)
    got = pointnet2_utils.stack_farthest_point_sample(T(xyz, cuda), T(cnt, cuda), [64, 32]).cpu().numpy()
    np.testing.assert_array_equal(got, O.stack_farthest_point_sample(xyz, cnt, [64, 32]))


@pytest.mark.parametrize("n,m,K,r", [(512, 512, 32, 0.3), (512, 64, 16, 0.1), (512, 100, 64, 1.0), (300, 77, 32, 0.5),
                                     (256, 256, 32, 0.2), (100, 30, 16, 0.7), (64, 64, 64, 0.3), (20, 7, 8, 0.4),
                                     (600, 50, 32, 0.3), (512, 33, 100, 2.0)])
def test_ball_query_uniform_layout(cuda, n, m, K, r):
    """fastpath.ball_query_uniform (register-resident warp-per-centroid kernel for n <= 512, the tiled kernel
    above that) against the oracle: wrapped clouds (duplicates), far-away centroids (empty balls), strided rows."""
    from multimodal_gar_b200 import fastpath
    Bp = 9
    rng = np.random.default_rng(n * 7 + K)
    c = cases.cloud_with_ties(n, Bp, n)
    c[::2] = rng.uniform(-1, 1, c[::2].shape).astype(np.float32)
    new_xyz = np.ascontiguousarray(c[:, rng.integers(0, n, m)] + rng.normal(0, 0.05, (Bp, m, 3)).astype(np.float32))
    new_xyz[:, ::5] += 50.0                                            # empty balls
    xyz = c.reshape(-1, 3)
    cnt, ncnt = np.full(Bp, n, np.int32), np.full(Bp, m, np.int32)
    want = O.ball_query_raw(r, K, xyz, cnt, new_xyz.reshape(-1, 3), ncnt) if hasattr(O, "ball_query_raw") else None
    wi, wm = O.ball_query(r, K, xyz, cnt, new_xyz.reshape(-1, 3), ncnt)
    c4 = np.concatenate([c, rng.random((Bp, n, 1)).astype(np.float32)], axis=2).reshape(-1, 4)
    for src in (T(xyz, cuda), T(c4, cuda)[:, 0:3]):
        raw = fastpath.ball_query_uniform(r, K, src, T(new_xyz.reshape(-1, 3), cuda), n, m).cpu().numpy()
        empty = raw[:, 0] == -1
        np.testing.assert_array_equal(empty, wm)
        assert (raw[empty, 1:] == 0).all()                             # fill = 1 writes every slot
        got = raw.copy()
        got[empty] = 0
        np.testing.assert_array_equal(got, wi)
        if want is not None:
            np.testing.assert_array_equal(raw, want)


@pytest.mark.parametrize("r,K", [(0.3, 16), (0.6, 32), (0.05, 8), (1.0, 64), (0.2, 1), (3.0, 100)])
def test_ball_query_and_grouping_ragged(cuda, r, K):
    xyz, cnt, new_xyz, ncnt, feats = cases.ragged_case()
    idx, mask = pointnet2_utils.ball_query(r, K, T(xyz, cuda), T(cnt, cuda), T(new_xyz, cuda), T(ncnt, cuda))
    want_i, want_m = O.ball_query(r, K, xyz, cnt, new_xyz, ncnt)
    np.testing.assert_array_equal(idx.cpu().numpy(), want_i)
    np.testing.assert_array_equal(mask.cpu().numpy(), want_m)
    out = pointnet2_utils.grouping_operation(T(feats, cuda), T(cnt, cuda), idx, T(ncnt, cuda))
    np.testing.assert_array_equal(out.cpu().numpy(), O.grouping_operation(feats, cnt, want_i, ncnt))


def test_ball_query_raw_sentinel(cuda):
    from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_stack import pointnet2_stack_cuda as ext
    xyz, cnt, new_xyz, ncnt, _ = cases.ragged_case()
    idx = torch.zeros((new_xyz.shape[0], 8), dtype=torch.int32, device=cuda)
    ext.ball_query_wrapper(len(cnt), new_xyz.shape[0], 0.05, 8, T(new_xyz, cuda), T(ncnt, cuda), T(xyz, cuda),
                           T(cnt, cuda), idx)
    want = O.ball_query_raw(0.05, 8, xyz, cnt, new_xyz, ncnt)
    assert (want[:, 0] == -1).any()
    np.testing.assert_array_equal(idx.cpu().numpy(), want)


def test_grouping_backward(cuda):
    xyz, cnt, new_xyz, ncnt, feats = cases.ragged_case()
    idx, _ = O.ball_query(0.6, 16, xyz, cnt, new_xyz, ncnt)
    f = T(feats, cuda).requires_grad_(True)
    out = pointnet2_utils.grouping_operation(f, T(cnt, cuda), T(idx, cuda), T(ncnt, cuda))
    rng = np.random.default_rng(0)
    g = rng.standard_normal(out.shape).astype(np.float32)
    out.backward(T(g, cuda))
    want = O.grouping_operation_grad(g, idx, ncnt, cnt, feats.shape[0])
    np.testing.assert_allclose(f.grad.cpu().numpy(), want, rtol=1e-5, atol=1e-5)   # float atomics: order differs


def test_query_and_group(cuda):
    xyz, cnt, new_xyz, ncnt, feats = cases.ragged_case()
    qg = pointnet2_utils.QueryAndGroup(0.3, 16, use_xyz=True)
    nf, idx = qg(T(xyz, cuda), T(cnt, cuda), T(new_xyz, cuda), T(ncnt, cuda), T(feats, cuda))
    want_nf, want_idx = O.query_and_group(0.3, 16, xyz, cnt, new_xyz, ncnt, feats)
    np.testing.assert_array_equal(idx.cpu().numpy(), want_idx)
    np.testing.assert_array_equal(nf.cpu().numpy(), want_nf)


def _feature_close(got, want):
    scale = max(1.0, float(np.abs(want).max()))
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-5 * scale)   # north_star: 1e-5 relative, fp32


@pytest.mark.parametrize("spec,K,r", [([2, 64, 128], 32, 0.6), ([2, 16], 16, 0.3), ([2, 32, 32, 64], 40, 0.6),
                                      ([2, 13, 7], 8, 0.05)])
def test_sa_module_eval_fused_mlp(cuda, spec, K, r):
    xyz, cnt, new_xyz, ncnt, feats = cases.ragged_case()
    full = [spec[0] + 3] + spec[1:]
    ref_mlp = O.make_sa_mlp(full, seed=4)
    mod = pointnet2_modules.StackSAModuleMSG(radii=[r], nsamples=[K], mlps=[list(spec)], use_xyz=True)
    mod.mlps[0].load_state_dict(ref_mlp.state_dict())
    mod = mod.to(cuda).eval()
    with torch.no_grad():
        _, got = mod(T(xyz, cuda), T(cnt, cuda), T(new_xyz, cuda), T(ncnt, cuda), T(feats, cuda))
    want = O.stack_sa_module_msg([ref_mlp], [r], [K], xyz, cnt, new_xyz, ncnt, feats)
    assert got.shape == want.shape
    _feature_close(got.cpu().numpy(), want)


def test_sa_module_two_scales_and_train_mode(cuda):
    xyz, cnt, new_xyz, ncnt, feats = cases.ragged_case()
    mlps = [O.make_sa_mlp([5, 16, 32], seed=1), O.make_sa_mlp([5, 8, 24], seed=2)]
    mod = pointnet2_modules.StackSAModuleMSG(radii=[0.3, 0.6], nsamples=[8, 16], mlps=[[2, 16, 32], [2, 8, 24]])
    for k in range(2):
        mod.mlps[k].load_state_dict(mlps[k].state_dict())
    mod = mod.to(cuda).eval()
    args = (T(xyz, cuda), T(cnt, cuda), T(new_xyz, cuda), T(ncnt, cuda), T(feats, cuda))
    with torch.no_grad():
        _, fused = mod(*args)
    want = O.stack_sa_module_msg(mlps, [0.3, 0.6], [8, 16], xyz, cnt, new_xyz, ncnt, feats)
    _feature_close(fused.cpu().numpy(), want)
    # with autograd on, the reference's own composition runs on the new grouping kernels: same numbers
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    _, composed = mod(*args)
    _feature_close(composed.detach().cpu().numpy(), want)
    composed.sum().backward()
    assert all(p.grad is not None for p in mod.parameters())


def test_sa_mlp_uniform_layout(cuda):
    rng = np.random.default_rng(5)
    Bp, n, m, K = 6, 256, 64, 16
    xyz = rng.uniform(-0.5, 0.5, (Bp * n, 3)).astype(np.float32)
    feats = rng.random((Bp * n, 1)).astype(np.float32)
    new_xyz = xyz.reshape(Bp, n, 3)[:, :m].reshape(-1, 3).copy()
    cnt, ncnt = np.full(Bp, n, np.int32), np.full(Bp, m, np.int32)
    mlp = O.make_sa_mlp([4, 64, 128], seed=9)
    idx, mask = O.ball_query(0.3, K, xyz, cnt, new_xyz, ncnt)
    ch, params = sa_mlp_fused.pack_shared_mlp(mlp)
    got = sa_mlp_fused.sa_mlp_forward(ch, params.to(cuda), K, T(xyz, cuda), None, T(new_xyz, cuda), None,
                                      T(feats, cuda), T(idx, cuda), uniform=(n, m))
    want = O.stack_sa_module_msg([mlp], [0.3], [K], xyz, cnt, new_xyz, ncnt, feats)
    _feature_close(got.cpu().numpy(), want)


@pytest.mark.parametrize("Bp,n,m,C_in,layout", [(40, 512, 512, 1, "uniform"), (7, 300, 129, 1, "ragged"),
                                                (5, 256, 64, 0, "ragged"), (3, 128, 50, 4, "uniform"),
                                                (6, 512, 100, 1, "strided")])
def test_sa_mlp_tensor_core_path(cuda, Bp, n, m, C_in, layout):
    """tcgen05 (3xTF32) path against the torch-CPU fp32 oracle AND the exact-fp32 CUDA kernel."""
    rng = np.random.default_rng(Bp * 1000 + n)
    K = 32
    pooled = rng.uniform(-0.5, 0.5, (Bp * n, 3 + max(C_in, 1))).astype(np.float32)
    pooled[:, :3] *= np.array([0.7, 0.7, 1.8], dtype=np.float32)
    xyz = np.ascontiguousarray(pooled[:, :3])
    feats = np.ascontiguousarray(pooled[:, 3:3 + C_in]) if C_in else None
    new_xyz = xyz.reshape(Bp, n, 3)[:, :m].reshape(-1, 3).copy()
    new_xyz[::37] += 9.0                                     # empty balls
    cnt, ncnt = np.full(Bp, n, np.int32), np.full(Bp, m, np.int32)
    mlp = O.make_sa_mlp([3 + C_in, 64, 128], seed=Bp)
    idx, mask = O.ball_query(0.3, K, xyz, cnt, new_xyz, ncnt)
    assert mask.any()
    want = O.stack_sa_module_msg([mlp], [0.3], [K], xyz, cnt, new_xyz, ncnt, feats)
    ch, params = sa_mlp_fused.pack_shared_mlp(mlp)
    assert sa_mlp_fused.tc_supported(ch, K, C_in)
    dpool = T(pooled, cuda)
    dxyz = dpool[:, :3] if layout == "strided" else T(xyz, cuda)
    kw = dict(uniform=(n, m)) if layout != "ragged" else {}
    cn, ncn = (None, None) if layout != "ragged" else (T(cnt, cuda), T(ncnt, cuda))
    dfe = None if feats is None else (dpool[:, 3:3 + C_in] if layout == "strided" else T(feats, cuda))
    zero = np.zeros(Bp, np.int32)
    zero[1] = 1
    gmax = torch.full((Bp, 128), -1.0, device=cuda)
    got = sa_mlp_fused.sa_mlp_forward(ch, params.to(cuda), K, dxyz, cn, T(new_xyz, cuda), ncn, dfe, T(idx, cuda),
                                      empty_mask=T(mask, cuda), params_host=params, impl='tc', group_max=gmax,
                                      group_size=m, group_zero=T(zero, cuda), **kw)
    _feature_close(got.cpu().numpy(), want)
    want_g = got.view(Bp, m, 128).max(dim=1).values
    want_g[1] = 0
    assert torch.equal(gmax, want_g)                         # fused per-cloud max == max of the rows it wrote
    only = sa_mlp_fused.sa_mlp_forward(ch, params.to(cuda), K, dxyz, cn, T(new_xyz, cuda), ncn, dfe, T(idx, cuda),
                                       empty_mask=T(mask, cuda), params_host=params, impl='tc',
                                       group_max=torch.empty_like(gmax), group_size=m, want_out=False, **kw)
    want_g[1] = got.view(Bp, m, 128)[1].max(dim=0).values
    assert torch.equal(only, want_g)
    if layout != "strided":
        ref32 = sa_mlp_fused.sa_mlp_forward(ch, params.to(cuda), K, dxyz, cn, T(new_xyz, cuda), ncn, dfe, T(idx, cuda),
                                            empty_mask=T(mask, cuda), impl='fp32', **kw)
        _feature_close(got.cpu().numpy(), ref32.cpu().numpy())


@pytest.mark.parametrize("n,m,K,radius", [(512, 512, 32, 0.3), (512, 100, 16, 0.6), (256, 256, 33, 0.4), (128, 300, 64, 0.25),
                                          (64, 64, 1, 0.5), (32, 7, 8, 0.2), (512, 1, 32, 0.3), (512, 512, 32, 1e-4)])
def test_fused_fps_ball_query_equals_the_two_kernels(cuda, n, m, K, radius):
    """One pass (sampling rounds double as ball-query distance tests) == fps_gather + ball_query_uniform, bit for bit:
    clouds with exact duplicates (early exit of the sampling, replicated rows), wrapped clouds, m > n, K > 32, a
    radius that only ever finds the centroid's own duplicates."""
    from multimodal_gar_b200 import fastpath
    b = 700
    c = cases.cloud_with_ties(n + m, b, n)
    rng = np.random.default_rng(n * 7 + m)
    c[::3] = rng.uniform(-1, 1, c[::3].shape).astype(np.float32)                 # every third cloud: no duplicates
    c[1::3, n // 4:] = c[1::3, :n - n // 4]                                     # wrapped clouds (k mod cnt)
    c4 = np.concatenate([c, rng.standard_normal((b, n, 1)).astype(np.float32)], axis=2)
    t4 = T(c4, cuda)
    xyz = t4[:, :, 0:3]
    fused = fastpath.fps_ball_query(xyz, m, radius, K)
    assert fused is not None
    idx, new_xyz = fastpath.fps_gather(xyz, m)
    ball = fastpath.ball_query_uniform(radius, K, t4.view(b * n, 4)[:, 0:3], new_xyz.view(b * m, 3), n, m)
    assert torch.equal(fused[0], idx) and torch.equal(fused[1], new_xyz)
    assert torch.equal(fused[2], ball)
    assert int((ball[:, 0] >= 0).sum()) == b * m
    pick = list(range(0, b, 100))                                                # and against the CPU oracle
    want = O.farthest_point_sample(np.ascontiguousarray(c[pick]), m)
    np.testing.assert_array_equal(fused[0].cpu().numpy()[pick], want)
    # contiguous (b,n,3) input and unsupported shapes
    f3 = fastpath.fps_ball_query(T(c, cuda), m, radius, K)
    assert all(torch.equal(a, b_) for a, b_ in zip(f3, fused))
    assert fastpath.fps_ball_query(T(c[:, :n - 1], cuda), m, radius, K) is None
    few = fastpath.fps_ball_query(T(c[:100], cuda), m, radius, K)
    if n >= 128:        # a CTA per cloud: any number of clouds
        assert torch.equal(few[0], idx[:100]) and torch.equal(few[2], ball[:100 * m])
    else:               # a warp per cloud: a throughput kernel, >= 592 clouds
        assert few is None
    # live_cnt = number of distinct samples: the indices behind it are 0, the ones before it distinct
    lc = fused[3].cpu().numpy()
    f_np = fused[0].cpu().numpy()
    for g in range(0, b, 23):
        assert 1 <= lc[g] <= m and (f_np[g, lc[g]:] == 0).all() and len(set(f_np[g, :lc[g]].tolist())) == lc[g]


@pytest.mark.parametrize("n,geom", [(512, "4,4"), (512, "2,8"), (512, "8,2"), (256, "4,2"), (256, "2,4"), (256, "8,1"),
                                    (128, "4,1"), (128, "2,2"), (128, "1,4"), (512, "warp")])
def test_fused_fps_ball_query_geometries_schedule_and_partial_rows(cuda, monkeypatch, n, geom):
    """Every CTA geometry of the fused kernel (and the warp-per-cloud kernel) writes the same bits; scheduling the clouds
    by weight changes nothing; with all_rows=False the rows of the distinct centroids (rounded up to four) are the same
    and the sampled indices / coordinates are complete."""
    from multimodal_gar_b200 import fastpath
    b, m, K, radius = 640, n, 32, 0.35
    c = cases.cloud_with_ties(n + 3, b, n)
    rng = np.random.default_rng(n)
    c[::2] = rng.uniform(-1, 1, c[::2].shape).astype(np.float32)
    for g in range(1, b, 4):                                                     # wrapped clouds with few distinct points
        k = int(rng.integers(1, n // 2))
        c[g] = c[g][np.arange(n) % k]
    x = T(c, cuda)
    monkeypatch.delenv("MGAR_FPS_BQ", raising=False)
    monkeypatch.delenv("MGAR_FPS_BQ_GEOM", raising=False)
    idx, new_xyz = fastpath.fps_gather(x, m)
    ball = fastpath.ball_query_uniform(radius, K, x.view(b * n, 3), new_xyz.view(b * m, 3), n, m)
    if geom == "warp":
        monkeypatch.setenv("MGAR_FPS_BQ", "warp")
    else:
        monkeypatch.setenv("MGAR_FPS_BQ_GEOM", geom)
    weight = T(rng.integers(0, 600, b).astype(np.int32), cuda)
    full = fastpath.fps_ball_query(x, m, radius, K, cloud_weight=weight)
    assert torch.equal(full[0], idx) and torch.equal(full[1], new_xyz) and torch.equal(full[2], ball)
    part = fastpath.fps_ball_query(x, m, radius, K, all_rows=False)
    assert torch.equal(part[0], idx) and torch.equal(part[1], new_xyz) and torch.equal(part[3], full[3])
    lc = full[3].cpu().numpy()
    rows = np.arange(m)[None, :] < ((lc + 3) // 4 * 4)[:, None]
    sel = torch.from_numpy(rows.reshape(-1)).to(cuda)
    assert torch.equal(part[2][sel], ball[sel])
    assert (lc[1::4] < n // 2).all() and (lc[0::2] == m).all()


@pytest.mark.parametrize("split", [False, True])
@pytest.mark.parametrize("m,K,radius", [(512, 32, 0.35), (512, 64, 0.6), (200, 16, 0.15), (512, 32, 5.0)])
def test_fused_fps_ball_query_periodic_clouds(cuda, monkeypatch, m, K, radius, split):
    """Wrapped clouds (slot k = point k mod p, what roipoint pooling produces) through the kernel that processes only the
    distinct blocks (fps_bq_dup.cuh): every period class - p < 32, p a multiple of 32, p = 511 / 512, periods whose copies
    straddle block boundaries, duplicated points INSIDE a period (ties between distinct slots) - with the true period, a
    multiple of it, a wrong hint (verified and ignored) and no hint; bit-exact against the two separate kernels.
    split: >= 1184 weighted clouds, the default launch - the clouds the hint calls long go to the four-warp kernel, the
    others to the periodic-cloud kernel beside it (a wrong hint only moves a cloud to the other kernel; "zero" sends every
    cloud to the periodic kernel, "double" most of them to the four-warp one)."""
    from multimodal_gar_b200 import fastpath
    monkeypatch.delenv("MGAR_FPS_BQ", raising=False)
    monkeypatch.delenv("MGAR_FPS_BQ_GEOM", raising=False)
    monkeypatch.delenv("MGAR_FPS_BQ_SPLIT", raising=False)
    if split:
        monkeypatch.delenv("MGAR_FPS_BQ_DUP", raising=False)
    else:
        monkeypatch.setenv("MGAR_FPS_BQ_DUP", "1")      # the periodic-cloud kernel alone
    n = 512
    periods = [1, 2, 5, 31, 32, 33, 63, 64, 65, 100, 145, 160, 200, 255, 256, 257, 300, 383, 384, 400, 479, 480, 481, 500, 511, 512]
    rng = np.random.default_rng(m + K)
    periods = periods + [int(v) for v in rng.integers(1, 513, (1300 if split else 640) - len(periods))]
    b = len(periods)
    c = np.zeros((b, n, 3), np.float32)
    for g, p in enumerate(periods):
        base = rng.uniform(-1, 1, (p, 3)).astype(np.float32)
        if g % 5 == 0 and p > 8:
            base[p // 2:p // 2 + 4] = base[:4]                                   # exact duplicates inside the period
        c[g] = base[np.arange(n) % p]
    c4 = np.concatenate([c, rng.standard_normal((b, n, 1)).astype(np.float32)], axis=2)
    t4 = T(c4, cuda)
    x = t4[:, :, 0:3]
    idx, new_xyz = fastpath.fps_gather(x, m)
    ball = fastpath.ball_query_uniform(radius, K, t4.view(b * n, 4)[:, 0:3], new_xyz.view(b * m, 3), n, m)
    true_p = T(np.array(periods, np.int32), cuda)
    hints = {"true": true_p, "double": torch.clamp(true_p * 2, max=600), "wrong": T(rng.integers(1, 513, b).astype(np.int32), cuda),
             "zero": torch.zeros_like(true_p)}
    for name, h in hints.items():
        n0 = _lib.launch_count()
        got = fastpath.fps_ball_query(x, m, radius, K, cloud_weight=h)
        assert _lib.launch_count() - n0 == (3 if split else 2), "order + %s" % ("two kernels side by side" if split else "one kernel")
        assert torch.equal(got[0], idx), name
        assert torch.equal(got[1], new_xyz), name
        assert torch.equal(got[2], ball), name
    none = fastpath.fps_ball_query(x, m, radius, K)
    assert torch.equal(none[2], ball) and torch.equal(none[3], got[3])
    part = fastpath.fps_ball_query(x, m, radius, K, all_rows=False, cloud_weight=true_p)
    lc = part[3].cpu().numpy()
    rows = np.arange(m)[None, :] < ((lc + 3) // 4 * 4)[:, None]
    sel = torch.from_numpy(rows.reshape(-1)).to(cuda)
    assert torch.equal(part[2][sel], ball[sel]) and torch.equal(part[0], idx)
