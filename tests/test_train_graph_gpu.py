"""-m gpu: configs[3]'s training step replayed from a CUDA graph (train_step.GraphedTrainStep): nothing in the step reads the
device back, so zero_grad + forward + backward + Adam are one capturable region; the graphed trajectory must be the eager
one (same kernels, same order: the same bits)."""
import pytest
import torch

from multimodal_gar_b200 import train_step as TS

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("kernels", ["", "tc"])
def test_graphed_training_step_follows_the_eager_trajectory(cuda, monkeypatch, kernels):
    """kernels = "tc": the statistics pass of the SA MLP on tcgen05 (MGAR_SA_TRAIN_KERNELS=tc) - its layer-1 parameters are
    read from the device, so it is capturable too."""
    if kernels:
        monkeypatch.setenv("MGAR_SA_TRAIN_KERNELS", kernels)
    else:
        monkeypatch.delenv("MGAR_SA_TRAIN_KERNELS", raising=False)

    def run(graphed, steps=4):
        torch.manual_seed(5)
        net = TS.LidarFusionNet().to(cuda).train()
        batch = TS.synthetic_batch(21, 1, N=32768, M=12, MNP=16, device=cuda)
        opt = torch.optim.Adam(net.parameters(), lr=1e-3, fused=True, capturable=graphed)
        if graphed:
            step = TS.GraphedTrainStep(net, opt, batch, warmup=2)
            losses = [float(step(batch)) for _ in range(steps)]
        else:
            for _ in range(2):
                TS.train_step(net, opt, batch)
            losses = [float(TS.train_step(net, opt, batch)) for _ in range(steps)]
        return losses, [p.detach().clone() for p in net.parameters()]

    l0, p0 = run(False)
    l1, p1 = run(True)
    assert l1 == l0, (l0, l1)
    assert l0[-1] < l0[0]
    for a, b in zip(p0, p1):
        assert torch.equal(a, b)
