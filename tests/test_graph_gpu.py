"""-m gpu: the LiDAR branch replayed from a CUDA graph (multimodal_gar_b200/graphs.py) - the library neither allocates nor
synchronises, so the whole step is capturable, the side-by-side sampling kernels (fork / join on the library's own stream)
included.  Replays on new inputs must be the bits of the eager step."""
import numpy as np
import pytest
import torch

from multimodal_gar_b200 import _lib, graphs, lidar_branch, synthetic

pytestmark = pytest.mark.gpu


def T(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


@pytest.mark.parametrize("B", [1, 32])          # 32 frames x 40 persons = 1280 clouds: the split sampling launch
def test_graph_replay_equals_the_eager_step(cuda, B):
    torch.manual_seed(3)
    enc = lidar_branch.PointPersonEncoder(1).to(cuda).eval()
    batches = []
    for seed in (11, 12, 13):
        pts, boxes = synthetic.make_batch(seed, B, N=65536, M=40)
        batches.append((T(pts[..., :3], cuda), T(pts[..., 3:4], cuda), T(boxes, cuda)))
    with torch.no_grad():
        eager = [enc(*b).clone() for b in batches]
    g = graphs.GraphedPersonEncoder(enc, *batches[0])
    n0 = _lib.launch_count()
    for b, want in zip(batches[1:] + batches[:1], eager[1:] + eager[:1]):
        got = g(*b)
        torch.cuda.synchronize()
        assert torch.equal(got, want)
    assert _lib.launch_count() == n0, "a replay must not go through the library's launch paths again"
    with pytest.raises(ValueError):
        g(batches[0][0][:, :100], batches[0][1][:, :100], batches[0][2])
    # a second graph of the same encoder (another resident batch) must leave the first one intact
    g2 = graphs.GraphedPersonEncoder(enc, *batches[1])
    enc.eval()                                   # a repeated eval() keeps the folded parameters the graphs point at
    for gg, b, want in ((g, batches[2], eager[2]), (g2, batches[0], eager[0]), (g, batches[1], eager[1])):
        got = gg(*b)
        torch.cuda.synchronize()
        assert torch.equal(got, want)
