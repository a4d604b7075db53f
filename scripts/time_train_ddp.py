"""configs[3] training step under DDP at N ranks (torchrun): ms per step with and without the gradient all-reduce, for a few
bucket sizes.  torchrun --nproc-per-node N scripts/time_train_ddp.py"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from multimodal_gar_b200 import train_step as TS  # noqa: E402


def timed(fn, steps, warm):
    for _ in range(warm):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / steps], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    batch = TS.synthetic_batch(300 + rank, 2, N=65536, M=80, MNP=100, device=dev)
    for cap in (25, 8, 4, 2, 1):
        torch.manual_seed(1994)
        net = TS.LidarFusionNet().to(dev).train()
        model = TS.make_ddp(net, dev, bucket_cap_mb=cap)
        opt = torch.optim.Adam(model.parameters(), lr=1e-4, fused=True)
        ms = timed(lambda: TS.train_step(model, opt, batch), 10, 3)

        def nosync():
            with model.no_sync():
                opt.zero_grad(set_to_none=True)
                model(batch).backward()
                opt.step()
        ms0 = timed(nosync, 10, 2)
        model(batch).backward()
        if rank == 0:
            print("bucket_cap_mb=%d world=%d ms_per_step=%.3f no_allreduce=%.3f exposed=%.3f" % (cap, dist.get_world_size(), ms, ms0, ms - ms0), flush=True)
        del model, net, opt
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
