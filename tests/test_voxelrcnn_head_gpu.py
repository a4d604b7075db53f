"""-m gpu: VoxelRCNNHead mirror (roi_grid_pool over three voxel scales + shared FC) on libmgar_lidar.so
against goldens produced by the REFERENCE class itself (tests/golden/make_ref_voxelrcnn_head_golden.py)."""
import importlib.util
import os

import numpy as np
import pytest
import torch

import cases
from multimodal_gar_b200.pcdet.models.roi_heads.voxelrcnn_head import SparseFeatures, VoxelRCNNHead

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _gen():
    spec = importlib.util.spec_from_file_location("make_ref_voxelrcnn_head_golden",
                                                  os.path.join(GOLD, "make_ref_voxelrcnn_head_golden.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    return gen


def _head(sc, g, cuda):
    head = VoxelRCNNHead(backbone_channels={"x_conv2": 8, "x_conv3": 12, "x_conv4": 16}, model_cfg=cases.voxelrcnn_head_cfg(),
                         point_cloud_range=[float(v) for v in sc["pc_range"]], voxel_size=[float(v) for v in sc["voxel_size"]])
    head.load_state_dict({k[5:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("head.")}, strict=True)
    return head.to(cuda)


def test_head_matches_reference_class_goldens(cuda):
    g = np.load(os.path.join(GOLD, "ref_voxelrcnn_head.npz"))
    sc = cases.voxelrcnn_scene()
    head = _head(sc, g, cuda)
    gen = _gen()
    tf32 = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False     # the goldens are fp32 (CPU)
    try:
        with torch.no_grad():
            head.eval()
            res = head(gen.batch_dict_of(sc, cuda, SparseFeatures))
            for key, tol in (("pooled_features", 1e-5), ("shared_feature", 2e-5)):
                want = g["eval." + key]
                assert tuple(res[key].shape) == want.shape
                np.testing.assert_allclose(res[key].cpu().numpy(), want, rtol=tol, atol=tol * float(np.abs(want).max()))
            head.train()
            want = g["train.pooled_features"]
            got = head.roi_grid_pool(gen.batch_dict_of(sc, cuda, SparseFeatures))
            np.testing.assert_allclose(got.cpu().numpy(), want, rtol=1e-4, atol=1e-4 * float(np.abs(want).max()))
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = tf32


def test_grid_points_and_frame_independence(cuda):
    g = np.load(os.path.join(GOLD, "ref_voxelrcnn_head.npz"))
    sc = cases.voxelrcnn_scene()
    head = _head(sc, g, cuda).eval()
    gen = _gen()
    boxes = torch.from_numpy(sc["gt_boxes"]).to(cuda)
    glob, local = head.get_global_grid_points_of_roi(boxes, grid_size=4)
    assert glob.shape == (8, 64, 3) and local.shape == (8, 64, 3)
    half = boxes.view(-1, 8)[:, 3:6].unsqueeze(1) / 2
    assert (local.abs() < half).all()                                    # cell centres lie strictly inside their box
    torch.testing.assert_close(local.mean(dim=1), torch.zeros(8, 3, device=cuda), atol=1e-6, rtol=0)
    # a frame's pooled features do not depend on its batch neighbour
    with torch.no_grad():
        both = head.roi_grid_pool(gen.batch_dict_of(sc, cuda, SparseFeatures))
        one = dict(sc)
        one["gt_boxes"] = sc["gt_boxes"][:1]
        for n in ("x_conv2", "x_conv3", "x_conv4"):
            keep = sc[n + ".indices"][:, 0] == 0
            one[n + ".indices"], one[n + ".features"] = sc[n + ".indices"][keep], sc[n + ".features"][keep]
        alone = head.roi_grid_pool(gen.batch_dict_of(one, cuda, SparseFeatures))
    assert torch.equal(both[:4], alone)
