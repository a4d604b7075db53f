"""-m gpu: THE step bench.py times - 64 configs[1] frames through PointPersonEncoder with its defaults - pinned end
to end: (a) it really runs the eight-kernel path (frame grid, box masks, pool-from-mask, cloud order, fused FPS + ball query
as two kernels side by side - four warps per long cloud, the periodic-cloud kernel for the wrapped ones - live-unit
list, tcgen05 MLP walking the distinct samples of the distinct centroids), (b) its person features are those of the
all-rows path (every ball row written, every centroid through the MLP: the same bits with the distinct-centroid walk, within
2e-6 with the distinct-sample walk whose 32-position blocks share their power-of-two operand scale differently), (c) every index / pooled value of that batch equals the
UNMODIFIED reference CUDA extensions (oracle/_ref) on the same GPU, bit for bit, and (d) sampled persons agree with the
reference's nn.Sequential on torch-CPU at 1e-5."""
import numpy as np
import pytest
import torch

import bench
import cases
from multimodal_gar_b200 import _lib
from multimodal_gar_b200.pcdet.ops.roipoint_pool3d import roipoint_pool3d_cuda as new_roipoint
from oracle import oracle as O

pytestmark = pytest.mark.gpu

F = 64


@pytest.fixture(scope="module")
def step(cuda):
    enc = bench.make_encoder().to(cuda)
    host, devb = bench.build_batches(F, 0, cuda)
    xyz, feat, bx = devb[1]                       # the permuted batch: frames 16.. are re-permuted copies
    with torch.no_grad():
        enc(xyz, feat, bx)                        # warm-up (lazy module state)
        torch.cuda.synchronize()
        n0 = _lib.launch_count()
        person = enc(xyz, feat, bx)
        torch.cuda.synchronize()
        launches = _lib.launch_count() - n0
        st = enc(xyz, feat, bx, return_stages=True)
        torch.cuda.synchronize()
    return dict(enc=enc, xyz=xyz, feat=feat, bx=bx, person=person, launches=launches, st=st)


def test_default_step_is_the_six_kernel_path_and_matches_the_all_rows_path(step):
    assert step["launches"] == 8, "expected frame grid, box masks, pool, cloud order, FPS+ball query (long | wrapped clouds), live units, MLP; got %d launches" % step["launches"]
    assert step["person"].shape == (F, bench.CFG["M"], 128)
    sa = step["st"]["sa_features"].view(F * bench.CFG["M"], bench.CFG["P"], 128)
    want = sa.max(dim=1).values * (step["st"]["empty"].view(-1, 1) == 0)
    assert torch.equal(step["st"]["person_features"].view(-1, 128), want)
    got = step["person"].view(-1, 128)
    scale = max(1.0, float(want.abs().max()))
    assert torch.allclose(got, want, rtol=2e-6, atol=2e-6 * scale), float((got - want).abs().max())
    assert torch.equal(got == 0, want == 0)
    # the distinct-centroid walk alone (whole centroids per TMEM lane quarter) keeps the bits
    enc = step["enc"]
    enc.skip_repeated_samples = False
    try:
        with torch.no_grad():
            tiles = enc(step["xyz"], step["feat"], step["bx"])
        torch.cuda.synchronize()
    finally:
        enc.skip_repeated_samples = True
    assert torch.equal(tiles.view(-1, 128), want)


def test_step_indices_equal_the_reference_extensions(step, cuda):
    roiaware, roipoint, pn2 = (O.load_ref(n) for n in ("roiaware_pool3d_cuda", "roipoint_pool3d_cuda", "pointnet2_stack_cuda"))
    if roiaware is None or roipoint is None or pn2 is None:
        pytest.skip("oracle/_ref not built")
    st, xyz, feat, bx = step["st"], step["xyz"], step["feat"], step["bx"]
    M, S, P, K, r = (bench.CFG[k] for k in ("M", "S", "P", "K", "radius"))
    N = xyz.shape[1]
    box_idx = torch.full((F, N), -1, dtype=torch.int32, device=cuda)
    roiaware.points_in_boxes_gpu(bx, xyz, box_idx)
    assert torch.equal(st["box_idx"], box_idx)
    pooled = torch.zeros((F, M, S, 4), device=cuda)
    empty = torch.zeros((F, M), dtype=torch.int32, device=cuda)
    roipoint.forward(xyz, bx, feat, pooled, empty)
    assert torch.equal(st["empty"], empty)
    # the step pools straight into the canonical frame; the same kernel family without the flag is the reference's rows
    mine = torch.empty_like(pooled)
    e2 = torch.empty_like(empty)
    new_roipoint.forward(xyz, bx, feat, mine, e2)
    assert torch.equal(mine, pooled)
    got = st["pooled"].view(F, M, S, 4).cpu().numpy()
    raw, boxes = pooled.cpu().numpy(), bx.cpu().numpy()
    for f in range(0, F, 7):
        want = O.canonical_transform(raw[f:f + 1], boxes[f:f + 1], empty[f:f + 1].cpu().numpy()).reshape(M, S, 4)
        live = empty[f].cpu().numpy() == 0
        cases.canonical_close(got[f][live], want[live], raw[f][live], boxes[f][live])
        assert (got[f][~live] == 0).all()
    # sampling and ball query of the reference on the step's own canonical clouds
    cloud = st["pooled"][:, :, 0:3].contiguous()
    b = F * M
    temp = torch.full((b, S), 1e10, device=cuda)
    fps = torch.zeros((b, P), dtype=torch.int32, device=cuda)
    pn2.farthest_point_sampling_wrapper(b, S, P, cloud, temp, fps)
    assert torch.equal(st["fps_idx"], fps)
    new_xyz = torch.gather(cloud, 1, fps.long().unsqueeze(-1).expand(-1, -1, 3)).reshape(b * P, 3).contiguous()
    assert torch.equal(st["new_xyz"], new_xyz)
    # distinct samples per person: behind live_cnt the sampling returns index 0 (what the MLP's walk relies on)
    lc = st["live_cnt"].cpu().numpy()
    f_np = fps.cpu().numpy()
    for g in range(0, b, 37):
        assert (f_np[g, lc[g]:] == 0).all() and len(set(f_np[g, :lc[g]].tolist())) == lc[g]
    # ball query of the reference (its kernel walks an O(batch) prefix per thread: the first 16 frames are enough)
    nb = 16 * M
    flat = cloud[:nb].reshape(nb * S, 3).contiguous()
    ball = torch.zeros((nb * P, K), dtype=torch.int32, device=cuda)
    cnt = torch.full((nb,), S, dtype=torch.int32, device=cuda)
    ncnt = torch.full((nb,), P, dtype=torch.int32, device=cuda)
    pn2.ball_query_wrapper(nb, nb * P, float(r), K, new_xyz[:nb * P].contiguous(), ncnt, flat, cnt, ball)
    assert torch.equal(st["ball_idx"][:nb * P], ball)


def test_sampled_persons_against_torch_cpu(step):
    st = step["st"]
    M, S, P, K, r = (bench.CFG[k] for k in ("M", "S", "P", "K", "radius"))
    empty = st["empty"].view(-1).cpu().numpy()
    live = np.flatnonzero(empty == 0)
    rng = np.random.default_rng(5)
    pick = sorted(set(rng.choice(live, 10, replace=False).tolist()) | {int(np.flatnonzero(empty)[0])} if empty.any() else
                  set(rng.choice(live, 10, replace=False).tolist()))
    pooled = st["pooled"].cpu().numpy()
    fps = st["fps_idx"].cpu().numpy()
    import copy
    mlp = copy.deepcopy(step["enc"].sa.mlps[0]).cpu()
    person = step["person"].view(-1, 128).cpu().numpy()
    for g in pick:
        cxyz = np.ascontiguousarray(pooled[g, :, :3])
        np.testing.assert_array_equal(O.farthest_point_sample(cxyz[None], P)[0], fps[g])
        new_xyz = cxyz[fps[g]]
        want = O.stack_sa_module_msg([mlp], [r], [K], cxyz, np.array([S], np.int32), new_xyz, np.array([P], np.int32),
                                     np.ascontiguousarray(pooled[g, :, 3:]))
        want = want.max(axis=0) * (empty[g] == 0)
        scale = max(1.0, float(np.abs(want).max()))
        np.testing.assert_allclose(person[g], want, rtol=1e-5, atol=1e-5 * scale)
