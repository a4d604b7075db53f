"""-m gpu: the new CUDA path against the UNMODIFIED reference CUDA extensions (oracle/_ref, built from
/root/reference sources by oracle/build_ref.py and shipped with the snapshot) on the same B200 and the
same inputs: every mask, index and gathered value bit-exact, no ambiguity allowance.  Also pins the
CPU oracle to the reference kernels at full BASELINE sizes."""
import numpy as np
import pytest
import torch

import cases
from multimodal_gar_b200 import synthetic
from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_stack import pointnet2_stack_cuda as new_pn2
from multimodal_gar_b200.pcdet.ops.roiaware_pool3d import roiaware_pool3d_cuda as new_roiaware
from multimodal_gar_b200.pcdet.ops.roipoint_pool3d import roipoint_pool3d_cuda as new_roipoint
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def T(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _ref(name):
    mod = O.load_ref(name)
    if mod is None:
        pytest.skip("oracle/_ref/%s.so not built" % name)
    return mod


@pytest.mark.parametrize("seed,N,M", [(1994, 65536, 40), (2023, 65536, 80), (0, 250000, 200), (1, 4099, 7)])
def test_points_in_boxes_gpu_bit_exact(cuda, seed, N, M):
    ref = _ref("roiaware_pool3d_cuda")
    pts, boxes = synthetic.make_batch(seed, 2, N=N, M=M)
    xyz, bx = T(pts[..., :3], cuda), T(boxes, cuda)
    a = torch.full((2, N), -1, dtype=torch.int32, device=cuda)
    b = torch.full((2, N), -1, dtype=torch.int32, device=cuda)
    ref.points_in_boxes_gpu(bx, xyz, a)
    new_roiaware.points_in_boxes_gpu(bx, xyz, b)
    assert (a >= 0).sum() > 100
    assert torch.equal(a, b)


def test_points_in_boxes_cpu_bit_exact(cuda):
    ref = _ref("roiaware_pool3d_cuda")
    pts, boxes = synthetic.make_frame(2, N=65536, M=40)
    xyz = torch.from_numpy(np.ascontiguousarray(pts[:, :3]))
    bx = torch.from_numpy(boxes)
    a = torch.zeros((40, 65536), dtype=torch.int32)
    b = torch.zeros((40, 65536), dtype=torch.int32)
    ref.points_in_boxes_cpu(bx, xyz, a)
    new_roiaware.points_in_boxes_cpu(bx, xyz, b)
    assert a.sum() > 100 and torch.equal(a, b)


@pytest.mark.parametrize("seed,N,M,S,extra", [(1994, 65536, 40, 512, 0.0), (2023, 65536, 80, 512, 0.2),
                                              (0, 250000, 200, 512, 0.0), (3, 4099, 7, 33, 1.0)])
def test_roipoint_pool_bit_exact(cuda, seed, N, M, S, extra):
    ref = _ref("roipoint_pool3d_cuda")
    pts, boxes = synthetic.make_batch(seed, 2, N=N, M=M, density_scale=4.0 if N > 100000 else 1.0)
    big = O.enlarge_box3d(boxes, (extra, extra, extra))
    xyz, feat, bx = T(pts[..., :3], cuda), T(pts[..., 3:4], cuda), T(big, cuda)
    pa = torch.zeros((2, M, S, 4), device=cuda)
    ea = torch.zeros((2, M), dtype=torch.int32, device=cuda)
    pb = torch.full((2, M, S, 4), 7.0, device=cuda)          # the new kernel writes every element
    eb = torch.full((2, M), 9, dtype=torch.int32, device=cuda)
    ref.forward(xyz, bx, feat, pa, ea)
    new_roipoint.forward(xyz, bx, feat, pb, eb)
    assert torch.equal(ea, eb)
    assert torch.equal(pa, pb)
    # and the CPU oracle agrees with the reference kernel wherever no pair is trig-ambiguous
    if N <= 65536:
        op, oe = O.roipoint_pool3d(pts[..., :3], pts[..., 3:4], boxes, (extra, extra, extra), S)
        for b in range(2):   # every box without a trig-ambiguous pair is compared
            _, box_ok = cases.unambiguous(O, pts[b, :, :3], big[b])
            assert box_ok.sum() >= M - 2
            np.testing.assert_array_equal(pa.cpu().numpy()[b][box_ok], op[b][box_ok])
            np.testing.assert_array_equal(ea.cpu().numpy()[b][box_ok], oe[b][box_ok])


@pytest.mark.parametrize("b,n,m,ties", [(40, 512, 512, True), (40, 512, 512, False), (8, 300, 128, True),
                                        (4, 1024, 256, True), (3, 2500, 300, True), (2, 20000, 128, False),
                                        (2, 700, 700, True), (5, 33, 20, True),
                                        # >= 592 clouds: the warp-per-cloud kernel (every ranks-per-lane variant)
                                        (640, 512, 512, True), (640, 512, 512, False), (600, 300, 128, True),
                                        (600, 200, 200, True), (700, 100, 64, False), (600, 64, 64, True),
                                        (600, 20, 20, True), (592, 1, 4, True)])
def test_fps_batch_bit_exact(cuda, b, n, m, ties):
    ref = _ref("pointnet2_stack_cuda")
    if ties:
        c = cases.cloud_with_ties(n + b, b, n)
    else:
        c = np.random.default_rng(n).uniform(-1, 1, (b, n, 3)).astype(np.float32)
    x = T(c, cuda)
    ta, tb = torch.full((b, n), 1e10, device=cuda), torch.full((b, n), 1e10, device=cuda)
    ia = torch.zeros((b, m), dtype=torch.int32, device=cuda)
    ib = torch.zeros((b, m), dtype=torch.int32, device=cuda)
    ref.farthest_point_sampling_wrapper(b, n, m, x, ta, ia)
    new_pn2.farthest_point_sampling_wrapper(b, n, m, x, tb, ib)
    assert torch.equal(ia, ib)
    assert torch.equal(ta, tb)
    if n <= 2500:
        k = min(b, 48)    # the scalar oracle is slow: a prefix of the clouds is enough
        np.testing.assert_array_equal(O.farthest_point_sample(c[:k], m), ia[:k].cpu().numpy())


def test_fps_stack_bit_exact(cuda):
    ref = _ref("pointnet2_stack_cuda")
    rng = np.random.default_rng(8)
    cnt = np.array([300, 1100, 77, 513, 5000], dtype=np.int32)
    npt = np.array([32, 700, 16, 513, 100], dtype=np.int32)
    xyz = np.concatenate([cases.cloud_with_ties(int(c), 1, int(c))[0] for c in cnt]) + \
        rng.integers(0, 2, (int(cnt.sum()), 1)).astype(np.float32)
    x = T(xyz, cuda)
    ta, tb = torch.full((xyz.shape[0],), 1e10, device=cuda), torch.full((xyz.shape[0],), 1e10, device=cuda)
    ia = torch.zeros((int(npt.sum()),), dtype=torch.int32, device=cuda)
    ib = torch.zeros_like(ia)
    ref.stack_farthest_point_sampling_wrapper(x, ta, T(cnt, cuda), ia, T(npt, cuda))
    new_pn2.stack_farthest_point_sampling_wrapper(x, tb, T(cnt, cuda), ib, T(npt, cuda))
    assert torch.equal(ia, ib)
    np.testing.assert_array_equal(O.stack_farthest_point_sample(xyz, cnt, npt), ia.cpu().numpy())


@pytest.mark.parametrize("r,K", [(0.1, 16), (0.3, 32), (0.5, 64), (1.0, 32), (0.02, 8)])
def test_ball_query_and_grouping_bit_exact_person_layout(cuda, r, K):
    """40 persons x 512 pooled points x 512 centroids: the configs[1] shapes."""
    ref = _ref("pointnet2_stack_cuda")
    rng = np.random.default_rng(int(r * 100) + K)
    Bp, n, m = 40, 512, 512
    xyz = (rng.uniform(-0.5, 0.5, (Bp * n, 3)) * np.array([0.7, 0.7, 1.8])).astype(np.float32)
    new_xyz = xyz.reshape(Bp, n, 3)[:, rng.permutation(n)[:m]].reshape(-1, 3).copy()
    new_xyz[::97] += 5.0                                     # some empty balls
    feats = rng.random((Bp * n, 4)).astype(np.float32)
    cnt, ncnt = T(np.full(Bp, n, np.int32), cuda), T(np.full(Bp, m, np.int32), cuda)
    ia = torch.zeros((Bp * m, K), dtype=torch.int32, device=cuda)
    ib = torch.zeros_like(ia)
    ref.ball_query_wrapper(Bp, Bp * m, r, K, T(new_xyz, cuda), ncnt, T(xyz, cuda), cnt, ia)
    new_pn2.ball_query_wrapper(Bp, Bp * m, r, K, T(new_xyz, cuda), ncnt, T(xyz, cuda), cnt, ib)
    assert (ia[:, 0] == -1).any()
    assert torch.equal(ia, ib)
    ia[ia[:, 0] == -1] = 0
    ga = torch.zeros((Bp * m, 4, K), device=cuda)
    gb = torch.full((Bp * m, 4, K), -3.0, device=cuda)
    ref.group_points_wrapper(Bp, Bp * m, 4, K, T(feats, cuda), cnt, ia, ncnt, ga)
    new_pn2.group_points_wrapper(Bp, Bp * m, 4, K, T(feats, cuda), cnt, ia, ncnt, gb)
    assert torch.equal(ga, gb)


def test_ball_query_ragged_bit_exact_and_oracle(cuda):
    ref = _ref("pointnet2_stack_cuda")
    xyz, cnt, new_xyz, ncnt, _ = cases.ragged_case()
    for r, K in ((0.3, 16), (0.6, 32), (0.05, 8)):
        ia = torch.zeros((new_xyz.shape[0], K), dtype=torch.int32, device=cuda)
        ib = torch.zeros_like(ia)
        ref.ball_query_wrapper(len(cnt), new_xyz.shape[0], r, K, T(new_xyz, cuda), T(ncnt, cuda), T(xyz, cuda),
                               T(cnt, cuda), ia)
        new_pn2.ball_query_wrapper(len(cnt), new_xyz.shape[0], r, K, T(new_xyz, cuda), T(ncnt, cuda), T(xyz, cuda),
                                   T(cnt, cuda), ib)
        assert torch.equal(ia, ib)
        np.testing.assert_array_equal(O.ball_query_raw(r, K, xyz, cnt, new_xyz, ncnt), ia.cpu().numpy())


def test_box_masks_crowded_cells_and_padded_slots_bit_exact(cuda):
    """More than 8 boxes over one grid cell (a crowd of overlapping persons; zero-padded proposal slots - degenerate boxes
    that all sit at the origin): the per-frame box grid marks the cell and only ITS points test every box.  First-hit
    indices and pooled rows against the reference extensions, bit for bit."""
    from multimodal_gar_b200 import fastpath
    roiaware, roipoint = _ref("roiaware_pool3d_cuda"), _ref("roipoint_pool3d_cuda")
    B, N, M = 3, 32768, 24
    pts, boxes = synthetic.make_batch(77, B, N=N, M=M)
    rng = np.random.default_rng(5)
    crowd = boxes[:, :12].copy()
    crowd[..., 0:2] = boxes[:, :1, 0:2] + rng.uniform(-0.25, 0.25, (B, 12, 2)).astype(np.float32)   # 12 persons in one cell
    boxes = np.concatenate([boxes, crowd, np.zeros((B, 20, 7), np.float32)], axis=1)                 # + 20 padded slots
    pts[:, :64, :3] = rng.uniform(-2e-5, 2e-5, (B, 64, 3)).astype(np.float32)                         # points at the origin
    pts[:, 64:128, :3] = 0.0
    M2 = boxes.shape[1]
    xyz, feat, bx = T(pts[..., :3], cuda), T(pts[..., 3:4], cuda), T(boxes, cuda)
    want_idx = torch.full((B, N), -1, dtype=torch.int32, device=cuda)
    roiaware.points_in_boxes_gpu(bx, xyz, want_idx)
    pa = torch.zeros((B, M2, 128, 4), device=cuda)
    ea = torch.zeros((B, M2), dtype=torch.int32, device=cuda)
    roipoint.forward(xyz, bx, feat, pa, ea)
    box_idx, mask = fastpath.box_masks(xyz, bx)
    pooled, empty, cnt = fastpath.roipoint_pool_from_masks(xyz, feat, bx, mask, 128, canonical=False, want_cnt=True)
    assert torch.equal(box_idx, want_idx)
    assert torch.equal(pooled, pa) and torch.equal(empty, ea)
    assert int((want_idx >= M).sum()) > 0                       # the crowd and / or the degenerate boxes really catch points
    assert torch.equal((cnt == 0).int(), ea)
