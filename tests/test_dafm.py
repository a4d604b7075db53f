"""CPU: the batched Distance Aware Fusion Module against golden vectors produced by the reference class itself
(tests/golden/make_ref_dafm_golden.py executes FusionAttention_mat cut out of gat_model.py:427-511 inside the
reference's per-frame loop :1396-1475).  fp32, tolerance 1e-5 relative to the largest feature (two attention +
LayerNorm stacks; De is computed as sqrt(|x|^2+|y|^2-2xy) here, by definition in fp64 there)."""
import os

import numpy as np
import torch

from multimodal_gar_b200 import dafm

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_dafm.npz"))


def _module():
    D = int(G["D"])
    fus = dafm.DistanceAwareFusion(D)
    for name, mod in (("m1.", fus.AttFusModule1), ("m2.", fus.AttFusModule2)):
        sd = {k[len(name):]: torch.from_numpy(G[k]) for k in G.files if k.startswith(name)}
        assert set(sd) == set(mod.state_dict()), "parameter names differ from the reference's"
        mod.load_state_dict(sd)
    return fus.eval()


def test_batched_dafm_matches_reference_per_frame_loop():
    fus = _module()
    counts = [int(c) for c in G["counts"]]
    MNP = G["rgb"].shape[1]
    valid = torch.arange(MNP).view(1, -1) < torch.tensor(counts).view(-1, 1)
    t = lambda k: torch.from_numpy(G[k])
    with torch.no_grad():
        got = fus(t("rgb"), t("lidar"), t("bboxes"), t("bboxes3d"), valid)
    want = t("fused")
    scale = float(want.abs().max())
    assert torch.allclose(got, want, rtol=1e-5, atol=1e-5 * scale), float((got - want).abs().max())
    assert (got[~valid] == 0).all()


def test_unbatched_call_is_the_reference_signature():
    fus = _module()
    n = int(G["counts"][0])
    t = lambda k: torch.from_numpy(G[k])
    R, L = t("rgb")[0, :n], t("lidar")[0, :n]
    Dg = dafm.generalized_box_iou(t("bboxes")[0, :n])
    from multimodal_gar_b200.lidar_backbone import pairwise_euclidean_distance
    De = pairwise_euclidean_distance(t("bboxes3d")[0, :n, :3])
    with torch.no_grad():
        Rp, Lp = fus.AttFusModule1(R, L, Dg, De)
        Rp, Lp = fus.AttFusModule2(Rp, Lp, Dg, De)
    want = t("fused")[0, :n]
    assert torch.allclose(torch.maximum(Rp, Lp), want, rtol=1e-5, atol=1e-5 * float(want.abs().max()))


def test_generalized_box_iou_matches_torchvision():
    import torchvision.ops as TO
    b = torch.from_numpy(G["bboxes"])[1]
    assert torch.allclose(dafm.generalized_box_iou(b), TO.generalized_box_iou(b, b), atol=1e-6)
