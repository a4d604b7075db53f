"""The drop-in contract: the REFERENCE's unmodified wrapper files - pointnet2_stack/pointnet2_utils.py,
pointnet2_modules.py, roiaware_pool3d_utils.py, roipoint_pool3d_utils.py (+ the box_utils / common_utils they import, with
a stub SharedArray) - imported over this repo's stand-in extension modules.  CPU part: they import and bind; -m gpu
part: the same Python file driven once over the stand-ins and once over the reference's compiled extensions gives the
same bits (indices, pooled rows, grouped features, gradients to float-atomic order), and the reference's own
StackSAModuleMSG class over the stand-ins matches this repo's fused module at 1e-5."""
import numpy as np
import pytest
import torch

import cases
import dropin
from multimodal_gar_b200 import _lib, synthetic

needs_ref = pytest.mark.skipif(not dropin.available(), reason="reference wrappers not available (oracle/_ref/pywrap)")


def T(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


@needs_ref
def test_reference_wrappers_import_over_the_stand_ins():
    m = dropin.load("new")
    pn2 = m["pointnet2_utils"]
    assert pn2.pointnet2.__name__.endswith("pointnet2_stack.pointnet2_stack_cuda") and "multimodal_gar_b200" in pn2.pointnet2.__name__
    for name in ("ball_query_wrapper", "group_points_wrapper", "group_points_grad_wrapper", "farthest_point_sampling_wrapper",
                 "stack_farthest_point_sampling_wrapper", "voxel_query_wrapper", "three_nn_wrapper", "three_interpolate_wrapper",
                 "three_interpolate_grad_wrapper", "vector_pool_wrapper", "vector_pool_grad_wrapper",
                 "query_stacked_local_neighbor_idxs_wrapper_stack", "query_three_nn_by_stacked_local_idxs_wrapper_stack"):
        assert callable(getattr(pn2.pointnet2, name)), name              # pointnet2_api.cpp:12-31, every exported name
    assert callable(m["roiaware_pool3d_utils"].roiaware_pool3d_cuda.points_in_boxes_gpu)
    assert callable(m["roipoint_pool3d_utils"].roipoint_pool3d_cuda.forward)
    for cls in ("StackSAModuleMSG", "StackPointnetFPModule"):
        assert hasattr(m["pointnet2_modules"], cls)
    # error behaviour: the reference prints and exit(-1)s on a CPU tensor (ball_query.cpp:14-26); the stand-in raises
    with pytest.raises(_lib.MgarError):
        m["roiaware_pool3d_utils"].points_in_boxes_gpu(torch.zeros(1, 8, 3), torch.zeros(1, 2, 7))


@needs_ref
@pytest.mark.gpu
def test_same_wrapper_two_back_ends_bit_exact(cuda):
    new, ref = dropin.load("new"), dropin.load("ref")
    pts, boxes = synthetic.make_batch(12, 2, N=32768, M=24)
    xyz, feat, bx = T(pts[..., :3], cuda), T(pts[..., 3:4], cuda), T(boxes, cuda)
    # roiaware_pool3d_utils.points_in_boxes_gpu / points_in_boxes_cpu (numpy in, as box_utils calls it)
    a = ref["roiaware_pool3d_utils"].points_in_boxes_gpu(xyz, bx)
    b = new["roiaware_pool3d_utils"].points_in_boxes_gpu(xyz, bx)
    assert (a >= 0).sum() > 100 and torch.equal(a, b)
    ca = ref["roiaware_pool3d_utils"].points_in_boxes_cpu(pts[0, :, :3], boxes[0])
    cb = new["roiaware_pool3d_utils"].points_in_boxes_cpu(pts[0, :, :3], boxes[0])
    ca, cb = np.asarray(ca), np.asarray(cb)                                  # numpy in -> numpy out (check_numpy_to_torch)
    assert ca.sum() > 100 and np.array_equal(ca, cb)
    # roipoint_pool3d_utils.RoIPointPool3d (enlarge_box3d from the reference's box_utils inside)
    pa, ea = ref["roipoint_pool3d_utils"].RoIPointPool3d(512, (0.2, 0.1, 0.0))(xyz, feat, bx)
    pb, eb = new["roipoint_pool3d_utils"].RoIPointPool3d(512, (0.2, 0.1, 0.0))(xyz, feat, bx)
    assert torch.equal(pa, pb) and torch.equal(ea, eb) and ea.sum() > 0
    # pointnet2_utils: FPS (batch + stack), ball_query, grouping (+ backward), QueryAndGroup
    cloud = pa.view(-1, 512, 4)[:, :, :3].contiguous()
    fa = ref["pointnet2_utils"].farthest_point_sample(cloud, 128)
    fb = new["pointnet2_utils"].farthest_point_sample(cloud, 128)
    assert torch.equal(fa, fb)
    x, cnt, q, qcnt, f = cases.ragged_case()
    X, C, Q, QC, F = T(x, cuda), T(cnt, cuda), T(q, cuda), T(qcnt, cuda), T(f, cuda)
    sa = ref["pointnet2_utils"].stack_farthest_point_sample(X, C, [16, 64, 8, 0, 33])
    sb = new["pointnet2_utils"].stack_farthest_point_sample(X, C, [16, 64, 8, 0, 33])
    assert torch.equal(sa, sb)
    ia, ma = ref["pointnet2_utils"].ball_query(0.4, 16, X, C, Q, QC)
    ib, mb = new["pointnet2_utils"].ball_query(0.4, 16, X, C, Q, QC)
    assert torch.equal(ia, ib) and torch.equal(ma, mb) and ma.any() and not ma.all()
    Fa, Fb = F.clone().requires_grad_(True), F.clone().requires_grad_(True)
    ga = ref["pointnet2_utils"].grouping_operation(Fa, C, ia, QC)
    gb = new["pointnet2_utils"].grouping_operation(Fb, C, ib, QC)
    assert torch.equal(ga, gb)
    w = torch.randn_like(ga)
    (ga * w).sum().backward()
    (gb * w).sum().backward()
    assert torch.allclose(Fa.grad, Fb.grad, rtol=1e-5, atol=1e-6)          # atomicAdd order
    qa = ref["pointnet2_utils"].QueryAndGroup(0.4, 16, use_xyz=True)(X, C, Q, QC, F)
    qb = new["pointnet2_utils"].QueryAndGroup(0.4, 16, use_xyz=True)(X, C, Q, QC, F)
    assert torch.equal(qa[0], qb[0]) and torch.equal(qa[1], qb[1])


@needs_ref
@pytest.mark.gpu
def test_reference_sa_module_class_over_the_stand_ins_matches_the_fused_module(cuda):
    """The reference's own StackSAModuleMSG (pointnet2_modules.py:30-112, cuDNN convolutions with TF32 off) running on
    the stand-in extension vs this repo's module with the same weights (tcgen05 kernel): 1e-5."""
    from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_stack import pointnet2_modules as mine_mod
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    new = dropin.load("new")
    torch.manual_seed(3)
    theirs = new["pointnet2_modules"].StackSAModuleMSG(radii=[0.3], nsamples=[32], mlps=[[1, 64, 128]], use_xyz=True,
                                                       pool_method="max_pool").to(cuda).eval()
    mine = mine_mod.StackSAModuleMSG(radii=[0.3], nsamples=[32], mlps=[[1, 64, 128]], use_xyz=True,
                                     pool_method="max_pool").to(cuda).eval()
    with torch.no_grad():
        for mod in theirs.modules():
            if isinstance(mod, torch.nn.BatchNorm2d):
                mod.running_mean.normal_(0, 0.1)
                mod.running_var.uniform_(0.8, 1.2)
                mod.weight.normal_(1, 0.2)
                mod.bias.normal_(0, 0.1)
    mine.load_state_dict(theirs.state_dict(), strict=True)                 # the reference's checkpoint keys load as they are
    rng = np.random.default_rng(0)
    cnt = np.array([700, 512, 33, 1200], np.int32)
    ncnt = np.array([128, 512, 40, 300], np.int32)
    X = T(rng.uniform(-1, 1, (int(cnt.sum()), 3)).astype(np.float32), cuda)
    Fe = T(rng.uniform(0, 1, (int(cnt.sum()), 1)).astype(np.float32), cuda)
    Q = T(rng.uniform(-1.1, 1.1, (int(ncnt.sum()), 3)).astype(np.float32), cuda)
    with torch.no_grad():
        _, want = theirs(X, T(cnt, cuda), Q, T(ncnt, cuda), Fe)
        _, got = mine(X, T(cnt, cuda), Q, T(ncnt, cuda), Fe)
    scale = float(want.abs().max())
    assert torch.allclose(got, want, rtol=1e-5, atol=1e-5 * scale), float((got - want).abs().max()) / scale
