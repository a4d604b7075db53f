import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..'))
import torch
from multimodal_gar_b200 import train_step as TS
dev = torch.device("cuda:0")
def run(graphed, steps=8):
    torch.manual_seed(1994)
    net = TS.LidarFusionNet().to(dev).train()
    batch = TS.synthetic_batch(300, 2, N=65536, M=80, MNP=100, device=dev)
    opt = torch.optim.Adam(net.parameters(), lr=1e-4, fused=True, capturable=graphed)
    losses = []
    if graphed:
        step = TS.GraphedTrainStep(net, opt, batch, warmup=3)
        f = lambda: step(batch).clone()
    else:
        for _ in range(3): TS.train_step(net, opt, batch)
        f = lambda: TS.train_step(net, opt, batch)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps): losses.append(f())
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps, [float(l) for l in losses]
ms0, l0 = run(False)
ms1, l1 = run(True)
print("eager %.3f ms" % ms0, l0[:3], l0[-1])
print("graph %.3f ms" % ms1, l1[:3], l1[-1])
