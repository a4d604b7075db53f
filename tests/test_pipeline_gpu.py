"""-m gpu: the composed LiDAR branch (PointPersonEncoder) at BASELINE sizes: the fused 5-kernel path
against the op-by-op composition through the drop-in pcdet.ops API, stage-wise against the oracle, and
size-independent properties (determinism, frame independence / sharding, empty persons)."""
import numpy as np
import pytest
import torch

import cases
from multimodal_gar_b200 import lidar_branch, synthetic
from multimodal_gar_b200.pcdet.ops.roiaware_pool3d import roiaware_pool3d_utils
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def T(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _inputs(seed, B, N, M, dev, **kw):
    pts, boxes = synthetic.make_batch(seed, B, N=N, M=M, **kw)
    return pts, boxes, T(pts[..., :3], dev), T(pts[..., 3:4], dev), T(boxes, dev)


def _enc(dev, **kw):
    torch.manual_seed(7)
    enc = lidar_branch.PointPersonEncoder(1, **kw).to(dev).eval()
    with torch.no_grad():                                   # non-trivial eval-mode BatchNorm
        for mod in enc.sa.mlps[0]:
            if isinstance(mod, torch.nn.BatchNorm2d):
                mod.running_mean.normal_(0, 0.1)
                mod.running_var.uniform_(0.8, 1.2)
                mod.weight.normal_(1, 0.1)
                mod.bias.normal_(0, 0.1)
    return enc


def test_fused_path_equals_op_by_op_composition_config2(cuda):
    pts, boxes, xyz, feat, bx = _inputs(1994, 2, 65536, 40, cuda)
    enc = _enc(cuda)
    with torch.no_grad():
        a = enc(xyz, feat, bx, return_stages=True)
        enc.sa.mlps[0].float()
        import os
        os.environ["MGAR_SA_MLP"] = "fp32"                  # force the exact-fp32 kernel and the composed path
        try:
            enc2 = _enc(cuda, nsample=32)
            enc2.load_state_dict(enc.state_dict())
            packed_supported = lidar_branch.sa_mlp_fused.tc_supported
            lidar_branch.sa_mlp_fused.tc_supported = lambda *a_, **k_: False
            b = enc2(xyz, feat, bx, return_stages=True)
        finally:
            lidar_branch.sa_mlp_fused.tc_supported = packed_supported
            del os.environ["MGAR_SA_MLP"]
    for key in ("box_idx", "empty", "fps_idx"):
        assert torch.equal(a[key], b[key]), key
    assert torch.equal(a["pooled"], b["pooled"])
    assert torch.equal(a["new_xyz"], b["new_xyz"])
    scale = float(b["sa_features"].abs().max())
    assert torch.allclose(a["sa_features"], b["sa_features"], rtol=1e-5, atol=1e-5 * scale)
    assert torch.allclose(a["person_features"], b["person_features"], rtol=1e-5, atol=1e-5 * scale)
    e = a["empty"].view(-1) > 0
    assert e.any() and (a["person_features"].view(-1, 128)[e] == 0).all()


def test_stagewise_against_oracle_config2_frame(cuda):
    pts, boxes, xyz, feat, bx = _inputs(2023, 1, 65536, 40, cuda)
    enc = _enc(cuda)
    with torch.no_grad():
        st = enc(xyz, feat, bx, return_stages=True)
    S, P, K, r, M = 512, 512, 32, 0.3, 40
    x_np, f_np = np.ascontiguousarray(pts[..., :3]), np.ascontiguousarray(pts[..., 3:4])
    # every point / box without a trig-ambiguous pair is compared (never the whole frame dropped)
    pt_ok, box_ok = cases.unambiguous(O, x_np[0], boxes[0])
    np.testing.assert_array_equal(st["box_idx"].cpu().numpy()[0][pt_ok], O.points_in_boxes_gpu(x_np, boxes)[0][pt_ok])
    pooled, empty = O.roipoint_pool3d(x_np, f_np, boxes, (0, 0, 0), S)
    assert box_ok.sum() >= M - 2
    np.testing.assert_array_equal(st["empty"].cpu().numpy()[0][box_ok], empty[0][box_ok])
    live = box_ok & (empty[0] == 0)
    cases.canonical_close(st["pooled"].cpu().numpy()[live], O.canonical_transform(pooled, boxes, empty)[live],
                          pooled[0][live], boxes[0][live])
    got_pooled = st["pooled"].cpu().numpy()
    cxyz = np.ascontiguousarray(got_pooled[..., :3])
    fps = O.farthest_point_sample(cxyz, P)
    np.testing.assert_array_equal(st["fps_idx"].cpu().numpy(), fps)
    new_xyz = np.take_along_axis(cxyz, fps[..., None].astype(np.int64), axis=1).reshape(-1, 3)
    np.testing.assert_array_equal(st["new_xyz"].cpu().numpy(), new_xyz)
    cnt, ncnt = np.full(M, S, np.int32), np.full(M, P, np.int32)
    want = O.stack_sa_module_msg([enc.sa.mlps[0].cpu()], [r], [K], cxyz.reshape(-1, 3), cnt, new_xyz, ncnt,
                                 np.ascontiguousarray(got_pooled[..., 3:]).reshape(-1, 1))
    scale = max(1.0, float(np.abs(want).max()))
    np.testing.assert_allclose(st["sa_features"].cpu().numpy(), want, rtol=1e-5, atol=1e-5 * scale)
    person = want.reshape(M, P, -1).max(axis=1) * (st["empty"].cpu().numpy().reshape(-1, 1) == 0)
    np.testing.assert_allclose(st["person_features"].cpu().numpy().reshape(M, -1), person, rtol=1e-5, atol=1e-5 * scale)


def test_frames_are_independent_and_deterministic(cuda):
    """Sharding property: a frame's result does not depend on which other frames share its batch."""
    pts, boxes, xyz, feat, bx = _inputs(3, 6, 65536, 40, cuda)
    enc = _enc(cuda)
    with torch.no_grad():
        full = enc(xyz, feat, bx)
        again = enc(xyz, feat, bx)
        assert torch.equal(full, again)
        for rank in range(2):
            lo, hi = lidar_branch.shard_frames(6, 2, rank)
            part = enc(xyz[lo:hi].contiguous(), feat[lo:hi].contiguous(), bx[lo:hi].contiguous())
            assert torch.equal(part, full[lo:hi])


def test_crowded_scene_stress_config5(cuda):
    """250k points x 200 boxes: pooled indices bit-exact against the oracle, per-person features finite."""
    pts, boxes, xyz, feat, bx = _inputs(5, 1, 250000, 200, cuda, density_scale=4.0)
    enc = _enc(cuda)
    with torch.no_grad():
        st = enc(xyz, feat, bx, return_stages=True)
    x_np, f_np = np.ascontiguousarray(pts[..., :3]), np.ascontiguousarray(pts[..., 3:4])
    _, box_ok = cases.unambiguous(O, x_np[0], boxes[0], max_amb=16)
    assert box_ok.sum() >= 190
    pooled, empty = O.roipoint_pool3d(x_np, f_np, boxes, (0, 0, 0), 512)
    np.testing.assert_array_equal(st["empty"].cpu().numpy()[0][box_ok], empty[0][box_ok])
    np.testing.assert_array_equal(st["pooled"].cpu().numpy()[box_ok][..., 3],
                                  O.canonical_transform(pooled, boxes, empty)[box_ok][..., 3])
    assert torch.isfinite(st["person_features"]).all()
    assert st["person_features"].shape == (1, 200, 128)


def test_point_lidar_backbone_output_contract(cuda):
    """a12: (B, MAX_NUM_PROPOSAL, 512) from points + zero-padded person boxes; padded rows = embedding(0)."""
    from multimodal_gar_b200 import lidar_backbone as LB
    pts, boxes, xyz, feat, bx = _inputs(11, 2, 16384, 12, cuda)
    MNP = 20
    padded = torch.zeros((2, MNP, 7), device=cuda)
    padded[:, :12] = bx
    torch.manual_seed(3)
    net = LB.PointLiDARBackbone().to(cuda).eval()
    with torch.no_grad():
        out = net(dict(points=T(pts, cuda), gt_bboxes3d=padded))
        person = net.encoder(xyz, feat, bx)
        assert out.shape == (2, MNP, 512)
        assert torch.allclose(out[:, :12], net.embedding(person), atol=1e-5)
        assert torch.allclose(out[:, 12:], net.embedding.bias.expand(2, MNP - 12, 512), atol=1e-6)
        pid = torch.full((2, MNP), -1, dtype=torch.int64, device=cuda)
        pid[:, :12] = torch.arange(12, device=cuda)
        d = LB.dafm_lidar_inputs(out, padded, pid)
    assert d["person_num"] == [12, 12] and d["De"].shape == (2, MNP, MNP)


@pytest.mark.parametrize("r,K", [(0.1, 16), (0.2, 32), (0.5, 64), (1.0, 16), (0.7, 32)])
def test_crowded_scene_radius_nsample_sweep_config5(cuda, r, K):
    """configs[4] sweep (radius 0.1-1.0, nsample 16-64) on a crowded frame: ball-query indices bit-exact and
    SA features within 1e-5 against the oracle on a sample of the 200 persons."""
    pts, boxes, xyz, feat, bx = _inputs(9, 1, 250000, 200, cuda, density_scale=4.0)
    enc = _enc(cuda, radius=r, nsample=K)
    with torch.no_grad():
        st = enc(xyz, feat, bx, return_stages=True)
    S, P = 512, 512
    pick = [0, 17, 63, 120, 199]
    pooled = st["pooled"].cpu().numpy()[pick]
    cxyz = np.ascontiguousarray(pooled[..., :3])
    fps = O.farthest_point_sample(cxyz, P)
    np.testing.assert_array_equal(st["fps_idx"].cpu().numpy()[pick], fps)
    new_xyz = np.take_along_axis(cxyz, fps[..., None].astype(np.int64), axis=1).reshape(-1, 3)
    cnt, ncnt = np.full(len(pick), S, np.int32), np.full(len(pick), P, np.int32)
    want = O.stack_sa_module_msg([enc.sa.mlps[0].cpu()], [r], [K], cxyz.reshape(-1, 3), cnt, new_xyz, ncnt,
                                 np.ascontiguousarray(pooled[..., 3:]).reshape(-1, 1))
    got = st["sa_features"].view(200, P, -1).cpu().numpy()[pick].reshape(-1, want.shape[1])
    scale = max(1.0, float(np.abs(want).max()))
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-5 * scale)


def test_host_frame_pipeline_matches_direct_calls(cuda):
    """The overlapped host-buffer front end returns, batch by batch, exactly what the encoder returns on device
    tensors (5 batches through 2 slots: slot reuse, stream hand-offs)."""
    from multimodal_gar_b200 import pipeline
    enc = _enc(cuda)
    pipe = pipeline.HostFramePipeline(enc, cuda, depth=2)
    hosts, outs, want = [], [], []
    for i in range(5):
        pts, boxes = synthetic.make_batch(100 + i, 2, N=16384, M=10)
        h = [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in (pts[..., :3], pts[..., 3:4], boxes)]
        hosts.append(h)
        outs.append(torch.empty((2, 10, 128)).pin_memory())
    for h, o in zip(hosts, outs):
        pipe.submit(*h, o)
    pipe.synchronize()
    with torch.no_grad():
        for h in hosts:
            want.append(enc(*[t.to(cuda) for t in h]).cpu())
    for o, w in zip(outs, want):
        assert torch.equal(o, w)


def test_config3_lidar_branch_plus_dafm_forward_is_frame_sharded(cuda):
    """configs[2] shape at reduced batch: 80 persons per frame (zero-padded to 100 proposals), full LiDAR branch +
    Linear 128->512 + two batched DAFM modules on synthetic RGB features; a frame's fused features do not depend on
    the other frames of its batch (what lets frames shard across GPUs with no collective)."""
    from multimodal_gar_b200 import dafm, lidar_backbone as LB
    B, M, MNP = 3, 80, 100
    pts, boxes, xyz, feat, bx = _inputs(21, B, 65536, M, cuda, density_scale=0.6)
    padded = torch.zeros((B, MNP, 7), device=cuda)
    padded[:, :M] = bx
    torch.manual_seed(9)
    net = LB.PointLiDARBackbone().to(cuda).eval()
    fus = dafm.DistanceAwareFusion(512).to(cuda).eval()
    rgb = torch.randn((B, MNP, 512), device=cuda)
    bb2 = torch.rand((B, MNP, 2), device=cuda) * 500
    bboxes = torch.cat([bb2, bb2 + 20 + 60 * torch.rand((B, MNP, 2), device=cuda)], dim=-1)
    pid = torch.full((B, MNP), -1, dtype=torch.int64, device=cuda)
    pid[:, :M] = torch.arange(M, device=cuda)
    pid[1, 50:] = -1                                              # frame 1 has only 50 persons
    with torch.no_grad():
        lidar = net(dict(points=T(pts, cuda), gt_bboxes3d=padded))
        d = LB.dafm_lidar_inputs(lidar, padded, pid)
        assert d["person_num"] == [80, 50, 80]
        fused = fus(rgb, d["feature"], bboxes, padded, d["valid"])
        assert fused.shape == (B, MNP, 512) and torch.isfinite(fused).all()
        assert (fused[1, 50:] == 0).all() and (fused[0, 80:] == 0).all()
        for b in range(B):
            one = net(dict(points=T(pts[b:b + 1], cuda), gt_bboxes3d=padded[b:b + 1]))
            d1 = LB.dafm_lidar_inputs(one, padded[b:b + 1], pid[b:b + 1])
            f1 = fus(rgb[b:b + 1], d1["feature"], bboxes[b:b + 1], padded[b:b + 1], d1["valid"])
            assert torch.equal(one, lidar[b:b + 1])
            assert torch.allclose(f1, fused[b:b + 1], rtol=1e-4, atol=1e-4)   # library GEMMs: batch size changes the kernel


def test_person_features_do_not_depend_on_storing_sa_features(cuda):
    """The (B*M*P, C) per-centroid SA features are an intermediate: with them kept on chip (default) the person
    features are the bits of max-over-centroids of the stored tensor with the distinct-centroid walk, and within 2e-6
    of them with the distinct-sample walk (the default; tests/test_live_units_gpu.py)."""
    _, _, xyz, feat, bx = _inputs(5, 3, 65536, 40, cuda)
    enc = _enc(cuda)
    with torch.no_grad():
        enc.keep_sa_features = False
        u = enc(xyz, feat, bx)
        enc.skip_repeated_samples = False
        a = enc(xyz, feat, bx)
        enc.keep_sa_features = True
        b = enc(xyz, feat, bx)
        st = enc(xyz, feat, bx, return_stages=True)
    assert torch.equal(a, b) and torch.equal(a, st["person_features"])
    want = st["sa_features"].view(3 * 40, 512, -1).max(dim=1).values
    want = want * (st["empty"].view(-1, 1) == 0)
    assert torch.equal(a.view(3 * 40, -1), want)
    scale = max(1.0, float(want.abs().max()))
    assert torch.allclose(u.view(3 * 40, -1), want, rtol=2e-6, atol=2e-6 * scale)


def test_folded_batchnorm_cache_follows_the_parameters(cuda):
    """The eval path folds BatchNorm into packed parameters and caches them: after load_state_dict() or an in-place edit
    (no .train()/.eval() in between) the next forward must use the NEW numbers."""
    pts, boxes, xyz, feat, bx = _inputs(4, 1, 16384, 8, cuda, density_scale=0.5)
    a, b = _enc(cuda), _enc(cuda)
    with torch.no_grad():
        for mod in b.sa.mlps[0]:
            if isinstance(mod, torch.nn.BatchNorm2d):
                mod.weight.normal_(0.3, 1.0)            # mixed-sign (and near-zero) scales: the sign-folded W2 rows
                mod.weight[0] = 0.0
                mod.running_mean.normal_(0, 0.3)
            if isinstance(mod, torch.nn.Conv2d):
                mod.weight.mul_(1.7)
        ya = a(xyz, feat, bx)                           # caches a's folded parameters
        want = b(xyz, feat, bx)
        assert not torch.equal(ya, want)
        a.load_state_dict(b.state_dict())               # in-place copy_: same storages, new versions
        assert torch.equal(a(xyz, feat, bx), want)
        a.sa.mlps[0][1].running_var.mul_(2.0)           # in-place edit of a buffer
        got = a(xyz, feat, bx)
        assert not torch.equal(got, want)
        c = _enc(cuda)
        c.load_state_dict(a.state_dict())
        assert torch.equal(c(xyz, feat, bx), got)
        # and the mixed-sign / zero BatchNorm scales agree with the exact-fp32 kernel and the oracle
        st = b(xyz, feat, bx, return_stages=True)
    cxyz = np.ascontiguousarray(st["pooled"].cpu().numpy()[..., :3])
    fps = st["fps_idx"].cpu().numpy()
    new_xyz = np.take_along_axis(cxyz, fps[..., None].astype(np.int64), axis=1).reshape(-1, 3)
    M, S, P = 8, 512, 512
    import copy
    want_sa = O.stack_sa_module_msg([copy.deepcopy(b.sa.mlps[0]).cpu()], [0.3], [32], cxyz.reshape(-1, 3), np.full(M, S, np.int32),
                                    new_xyz, np.full(M, P, np.int32),
                                    np.ascontiguousarray(st["pooled"].cpu().numpy()[..., 3:]).reshape(-1, 1))
    scale = max(1.0, float(np.abs(want_sa).max()))
    np.testing.assert_allclose(st["sa_features"].cpu().numpy(), want_sa, rtol=1e-5, atol=1e-5 * scale)


def test_single_process_two_devices(cuda):
    """One process driving two GPUs (the reference's own launcher is nn.DataParallel, train_func.py:512): shared-memory
    opt-ins and scratch are per device / caller-owned, so the second device must give the first one's bits."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs in one process")
    from multimodal_gar_b200 import dafm
    pts, boxes = synthetic.make_batch(9, 2, N=32768, M=24)
    outs = []
    for d in (0, 1):
        dev = torch.device("cuda", d)
        enc = _enc(dev)
        torch.manual_seed(4)
        fus = dafm.FusionAttentionMat(512, 512).to(dev).eval()
        R = torch.randn((2, 24, 512), generator=torch.Generator().manual_seed(1)).to(dev)
        L = torch.randn((2, 24, 512), generator=torch.Generator().manual_seed(2)).to(dev)
        De = torch.cdist(T(boxes[..., :3], dev), T(boxes[..., :3], dev))
        with torch.no_grad():
            st = enc(T(pts[..., :3], dev), T(pts[..., 3:4], dev), T(boxes, dev), return_stages=True)
            m = roiaware_pool3d_utils.points_in_boxes_cpu(torch.from_numpy(np.ascontiguousarray(pts[0, :, :3])), torch.from_numpy(boxes[0]))
            f = fus(R, L, None, De)
        outs.append((st["box_idx"].cpu(), st["pooled"].cpu(), st["fps_idx"].cpu(), st["person_features"].cpu(), m.cpu(),
                     f[0].cpu(), f[1].cpu()))
    for a, b in zip(*outs):
        assert torch.equal(a, b)
    # the default batch path on the second device too: 1280 clouds -> the two sampling kernels side by side on the
    # library's per-device side stream, the live-unit list, the tcgen05 MLP
    pts, boxes = synthetic.make_batch(10, 32, N=32768, M=40)
    big = []
    for d in (0, 1):
        dev = torch.device("cuda", d)
        enc = _enc(dev)
        with torch.no_grad():
            big.append(enc(T(pts[..., :3], dev), T(pts[..., 3:4], dev), T(boxes, dev)).cpu())
    assert torch.equal(big[0], big[1]) and big[0].abs().sum() > 0
