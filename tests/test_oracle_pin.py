"""Pins the CPU oracle (oracle/mgar_oracle.c) to the reference itself.

(i)  orc_points_in_boxes_mask == the reference's points_in_boxes_cpu, compiled unmodified into
     oracle/_ref (present in the build container and shipped to the GPU box), and == the committed
     fixture tests/golden/ref_cpu.npz generated from it (tests/golden/make_ref_cpu_golden.py).
(ii) every other oracle function == tests/golden/ref_gpu.npz, the outputs of the reference's CUDA
     extensions on a B200 (tests/golden/make_ref_gpu_golden.py).
(iii) the FPS oracle == an independent pure-Python thread-by-thread simulation of the reference's
     block (sampling_gpu.cu:25-140), small cases.
"""
import os

import numpy as np
import pytest

import cases
from oracle import oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _ref_cpu_mask(points, boxes):
    import torch
    mod = O.load_ref("roiaware_pool3d_cuda")
    if mod is None:
        pytest.skip("oracle/_ref/roiaware_pool3d_cuda.so not built")
    out = torch.zeros((boxes.shape[0], points.shape[0]), dtype=torch.int32)
    mod.points_in_boxes_cpu(torch.from_numpy(boxes).contiguous(), torch.from_numpy(points).contiguous(), out)
    return out.numpy()


@pytest.mark.parametrize("seed", [1994, 2023, 0])
def test_mask_oracle_equals_reference_cpu_function(seed):
    from multimodal_gar_b200 import synthetic
    pts, boxes = synthetic.make_frame(seed, N=16384, M=40)
    xyz = np.ascontiguousarray(pts[:, :3])
    want = _ref_cpu_mask(xyz, boxes)
    got = O.points_in_boxes_cpu(xyz, boxes)
    assert want.sum() > 100
    np.testing.assert_array_equal(got, want)


def test_mask_oracle_equals_reference_on_face_points():
    xyz, _, boxes = cases.frame_case()
    for b in range(xyz.shape[0]):
        np.testing.assert_array_equal(O.points_in_boxes_cpu(xyz[b], boxes[b]), _ref_cpu_mask(xyz[b], boxes[b]))


def test_mask_oracle_equals_committed_cpu_fixture():
    path = os.path.join(GOLD, "ref_cpu.npz")
    assert os.path.exists(path), "tests/golden/ref_cpu.npz missing (make_ref_cpu_golden.py)"
    g = np.load(path)
    got = O.points_in_boxes_cpu(g["xyz"], g["boxes"])
    np.testing.assert_array_equal(np.packbits(got.astype(np.uint8), axis=1), g["mask_bits"])
    assert int(got.sum()) == int(g["hits"])


# ---- (iii) literal simulation of the reference FPS block -------------------------------------------
def _fps_block_sim(xyz, m, bs):
    """sampling_gpu.cu:25-140: bs threads, thread t scans k=t,t+bs,.. (first strict max), then the
    shared-memory tree for strides bs/2..1 with strict '>' (__update :16-21)."""
    f = np.float32
    n = xyz.shape[0]
    temp = np.full(n, 1e10, dtype=np.float32)
    idx = np.zeros(m, dtype=np.int32)
    old = 0
    for j in range(1, m):
        dists = np.full(bs, -1, dtype=np.float32)
        dists_i = np.zeros(bs, dtype=np.int64)
        x1, y1, z1 = xyz[old]
        for t in range(bs):
            best, besti = f(-1), 0
            for k in range(t, n, bs):
                dx, dy, dz = f(xyz[k, 0] - x1), f(xyz[k, 1] - y1), f(xyz[k, 2] - z1)
                # fma(dz,dz, fma(dx,dx, dy*dy)) evaluated exactly in float64 then rounded once per fma
                d = f(np.float64(dx) * np.float64(dx) + np.float64(f(dy * dy)))
                d = f(np.float64(dz) * np.float64(dz) + np.float64(d))
                d2 = min(d, temp[k])
                temp[k] = d2
                if d2 > best:
                    best, besti = d2, k
            dists[t], dists_i[t] = best, besti
        s = bs // 2
        while s >= 1:
            for t in range(s):
                if dists[t + s] > dists[t]:
                    dists[t], dists_i[t] = dists[t + s], dists_i[t + s]
            s //= 2
        old = int(dists_i[0])
        idx[j] = old
    return idx


@pytest.mark.parametrize("n,m", [(1, 1), (5, 5), (37, 20), (64, 64), (100, 40), (130, 33)])
def test_fps_oracle_equals_block_simulation(n, m):
    c = cases.cloud_with_ties(n + 11, 1, n)[0]
    bs = O.opt_n_threads(n)
    assert bs == min(1024, 1 << (int(n).bit_length() - 1))
    np.testing.assert_array_equal(O.farthest_point_sample(c[None], m)[0], _fps_block_sim(c, m, bs))


def test_opt_n_threads_is_floor_pow2_capped():
    for n in list(range(1, 70)) + [255, 256, 257, 511, 512, 513, 1023, 1024, 1025, 4096, 65536, 250000]:
        assert O.opt_n_threads(n) == min(1024, 1 << (n.bit_length() - 1))


# ---- (ii) the reference CUDA extensions' outputs on a B200 ----------------------------------------
def _gold_gpu():
    path = os.path.join(GOLD, "ref_gpu.npz")
    if not os.path.exists(path):
        pytest.skip("tests/golden/ref_gpu.npz not generated yet (needs one GPU run)")
    return np.load(path)


def test_oracle_equals_reference_gpu_in_box_and_pooling():
    g = _gold_gpu()
    xyz, feat, boxes = g["f_xyz"], g["f_feat"], g["f_boxes"]
    got = O.points_in_boxes_gpu(xyz, boxes)
    amb = np.stack([O.in_box_ambiguous(xyz[b], boxes[b]).any(axis=0) for b in range(xyz.shape[0])])
    diff = (got != g["pib_gpu"]) & ~amb           # libdevice vs glibc trig may flip 1-ulp-ambiguous pairs only
    assert not diff.any()
    assert (got != g["pib_gpu"]).sum() <= amb.sum()
    np.testing.assert_array_equal(O.points_in_boxes_cpu(xyz[0], boxes[0]), g["pib_cpu_frame0"])
    for tag, extra, S in (("e0", (0.0, 0.0, 0.0), 128), ("e2", (0.2, 0.2, 0.2), 64)):
        pooled, empty = O.roipoint_pool3d(xyz, feat, boxes, extra, S)
        np.testing.assert_array_equal(empty, g["pool_empty_%s" % tag])
        np.testing.assert_array_equal(pooled, g["pool_%s" % tag])


def test_oracle_equals_reference_gpu_fps():
    g = _gold_gpu()
    for n, m in ((128, 64), (200, 64), (512, 512), (1500, 96), (2500, 40)):
        c = cases.cloud_with_ties(n, 3, n)
        idx, temp = O.farthest_point_sample(c, m, return_temp=True)
        np.testing.assert_array_equal(idx, g["fps_%d" % n])
        np.testing.assert_array_equal(temp, g["fps_temp_%d" % n])
    pos = g["r_cnt"] > 0
    np.testing.assert_array_equal(O.stack_farthest_point_sample(g["r_xyz"], g["r_cnt"][pos], g["sfps_npoint"]),
                                  g["sfps"])


def test_oracle_equals_reference_gpu_ball_query_and_grouping():
    g = _gold_gpu()
    for r, K in ((0.3, 16), (0.6, 32), (0.05, 8)):
        raw = O.ball_query_raw(r, K, g["r_xyz"], g["r_cnt"], g["r_new"], g["r_ncnt"])
        np.testing.assert_array_equal(raw, g["bq_%g_%d" % (r, K)])
        fixed = raw.copy()
        fixed[fixed[:, 0] == -1] = 0
        np.testing.assert_array_equal(O.grouping_operation(g["r_feat"], g["r_cnt"], fixed, g["r_ncnt"]),
                                      g["grp_%g_%d" % (r, K)])


def test_oracle_equals_reference_gpu_roiaware_pool3d():
    """RoIAwarePool3d: outputs of the unmodified reference kernels on a B200
    (tests/golden/make_ref_roiaware_golden.py) against the C restatement - forward bit-exact, backward to the
    order of the reference's float atomics."""
    path = os.path.join(GOLD, "ref_roiaware.npz")
    if not os.path.exists(path):
        pytest.skip("tests/golden/ref_roiaware.npz not generated yet (needs one GPU run)")
    g = np.load(path)
    size, max_pts = tuple(int(v) for v in g["size"]), int(g["max_pts"])
    for method in ("max", "avg"):
        pooled, argmax, lists = O.roiaware_pool3d(g["boxes"], g["pts"], g["feat"], size, max_pts, method)
        np.testing.assert_array_equal(lists, g[method + "_lists"])
        np.testing.assert_array_equal(argmax, g[method + "_argmax"])
        np.testing.assert_array_equal(pooled, g[method + "_pooled"])
        assert lists[..., 0].max() == max_pts - 1 and (lists[..., 0] == 0).any()      # capped and empty voxels both present
        gin = O.roiaware_pool3d_backward(lists, argmax, g[method + "_grad_out"], g["pts"].shape[0], method)
        np.testing.assert_allclose(gin, g[method + "_grad_in"], rtol=1e-5, atol=1e-6)


def test_oracle_equals_reference_gpu_voxel_query():
    """Voxel query: rows written by the unmodified reference kernel on a B200
    (tests/golden/make_ref_voxel_query_golden.py) against the C restatement, bit-exact."""
    path = os.path.join(GOLD, "ref_voxel_query.npz")
    if not os.path.exists(path):
        pytest.skip("tests/golden/ref_voxel_query.npz not generated yet (needs one GPU run)")
    g = np.load(path)
    src = open(os.path.join(GOLD, "make_ref_voxel_query_golden.py")).read()
    scenes = eval(src[src.index("SCENES = ") + 9:src.index("\n\n\ndef main")])      # the scene list of the generating script
    for i, case in enumerate(scenes):
        c = dict(case)
        rng, radius, nsample = c.pop("rng"), c.pop("radius"), c.pop("nsample")
        sc = cases.voxel_scene(C=1, **c)
        got = O.voxel_query_raw(rng, radius, nsample, sc["xyz"], sc["new_xyz"], sc["new_coords"], sc["v2p"])
        np.testing.assert_array_equal(got, g["idx_%d" % i])
        assert (got[:, 0] == -1).any() and (got[:, 0] >= 0).any()


def test_oracle_equals_reference_gpu_batch_ops():
    """pointnet2_batch surface: outputs of the unmodified reference kernels on a B200
    (tests/golden/make_ref_batch_golden.py) against the C restatement, bit-exact."""
    path = os.path.join(GOLD, "ref_batch_ops.npz")
    if not os.path.exists(path):
        pytest.skip("tests/golden/ref_batch_ops.npz not generated yet (needs one GPU run)")
    g = np.load(path)
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_ref_batch_golden", os.path.join(GOLD, "make_ref_batch_golden.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    xyz, new_xyz, feats, weight = gen.inputs()
    bq = O.batch_ball_query(0.35, 16, xyz, new_xyz)
    np.testing.assert_array_equal(bq, g["bq"])
    assert (bq[:, :2] == 0).all() and (bq[:, 2:, 1] > 0).any()
    _, idx, d2 = O.three_nn(xyz, new_xyz)
    np.testing.assert_array_equal(idx, g["nn_idx"])
    np.testing.assert_array_equal(d2, g["d2"])
    np.testing.assert_array_equal(O.three_interpolate(feats, idx, weight), g["interp"])


def test_oracle_equals_reference_gpu_stack_interp():
    """stack three_nn / three_interpolate: outputs of the unmodified reference kernels on a B200
    (tests/golden/make_ref_stack_interp_golden.py) against the C restatement, bit-exact."""
    path = os.path.join(GOLD, "ref_stack_interp.npz")
    if not os.path.exists(path):
        pytest.skip("tests/golden/ref_stack_interp.npz not generated yet (needs one GPU run)")
    g = np.load(path)
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_ref_stack_interp_golden",
                                                  os.path.join(GOLD, "make_ref_stack_interp_golden.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    unknown, ucnt, known, kcnt, feats, weight = gen.inputs()
    _, idx, d2 = O.stack_three_nn(unknown, ucnt, known, kcnt)
    np.testing.assert_array_equal(idx, g["nn_idx"])
    np.testing.assert_array_equal(d2, g["d2"])
    assert np.isinf(d2).any() and (idx[ucnt[0] + ucnt[1]:ucnt[0] + ucnt[1] + ucnt[2], 2] == kcnt[0] + kcnt[1]).all()
    np.testing.assert_array_equal(O.stack_three_interpolate(feats, np.clip(idx, 0, known.shape[0] - 1), weight), g["interp"])
