"""Where the configs[3] training step spends its time: torch.profiler over a few steps (CUDA kernels + CPU ops)."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from multimodal_gar_b200 import train_step as TS  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    frames = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    torch.manual_seed(1994)
    net = TS.LidarFusionNet().to(dev).train()
    batch = TS.synthetic_batch(300, frames, N=65536, M=80, MNP=100, device=dev)
    opt = torch.optim.Adam(net.parameters(), lr=1e-4)
    for _ in range(3):
        TS.train_step(net, opt, batch)
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        for _ in range(3):
            TS.train_step(net, opt, batch)
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=70))
    print(prof.key_averages().table(sort_by="cpu_time_total", row_limit=15, max_name_column_width=70))


if __name__ == "__main__":
    main()
