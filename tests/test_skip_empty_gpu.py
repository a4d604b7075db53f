"""-m gpu: the tensor-core SA MLP walks only the persons that pooled points (device-side live list, TcArgs::live_groups).
Person features must be the same bits with the skip on (PointPersonEncoder(skip_empty_persons=True), what
the default) and off, for mixed, all-empty, single-live and
zero-padded frames."""
import numpy as np
import pytest
import torch

from multimodal_gar_b200 import _lib, lidar_branch, synthetic

pytestmark = pytest.mark.gpu


def _enc(dev):
    torch.manual_seed(7)
    enc = lidar_branch.PointPersonEncoder(1).to(dev).eval()
    enc.skip_repeated_samples = False        # whole centroids per TMEM lane quarter: the walks below keep the bits
    with torch.no_grad():
        for mod in enc.sa.mlps[0]:
            if isinstance(mod, torch.nn.BatchNorm2d):
                mod.running_mean.normal_(0, 0.1)
                mod.running_var.uniform_(0.8, 1.2)
                mod.weight.normal_(1, 0.1)
                mod.bias.normal_(0, 0.1)
    return enc


def _run(enc, pts, boxes, dev, monkeypatch, skip):
    monkeypatch.delenv("MGAR_TC_SKIP_EMPTY", raising=False)
    enc.skip_empty_persons = skip
    xyz = torch.from_numpy(np.ascontiguousarray(pts[..., :3])).to(dev)
    feat = torch.from_numpy(np.ascontiguousarray(pts[..., 3:4])).to(dev)
    n0 = _lib.launch_count()
    with torch.no_grad():
        st = enc(xyz, feat, torch.from_numpy(boxes).to(dev), return_stages=False)
    torch.cuda.synchronize()
    out = st["person_features"] if isinstance(st, dict) else st
    return out, _lib.launch_count() - n0


@pytest.mark.parametrize("case", ["all_empty", "one_live", "mixed", "padded"])
def test_person_features_identical_with_and_without_the_skip(cuda, monkeypatch, case):
    B, M = (2, 8) if case in ("all_empty", "one_live") else (3, 40)
    pts, boxes = synthetic.make_batch(31, B, N=32768, M=M, density_scale=0.6)
    if case == "all_empty":
        boxes[..., :2] += 500.0
    elif case == "one_live":
        keep = boxes[1, 3].copy()
        boxes[..., :2] += 500.0
        boxes[1, 3] = keep
    elif case == "padded":
        boxes = np.concatenate([boxes, np.zeros((B, 24, 7), np.float32)], axis=1)
    enc = _enc(cuda)
    on, n_on = _run(enc, pts, boxes, cuda, monkeypatch, True)
    off, n_off = _run(enc, pts, boxes, cuda, monkeypatch, False)
    assert n_on == n_off + 1, "the live-person compaction kernel did not run"
    assert torch.equal(on, off)
    flat = on.reshape(-1, on.shape[-1])
    if case == "all_empty":
        assert (flat == 0).all()
    if case == "one_live":
        assert int((flat.abs().sum(dim=1) > 0).sum()) == 1
    if case == "padded":
        assert (on[:, 40:] == 0).all() and (on[:, :40].abs().sum() > 0)


def test_backbone_for_padded_proposals_uses_the_skip(cuda):
    from multimodal_gar_b200 import lidar_backbone
    assert lidar_backbone.PointLiDARBackbone().encoder.skip_empty_persons
    assert lidar_branch.PointPersonEncoder(1).skip_empty_persons                # default: on
