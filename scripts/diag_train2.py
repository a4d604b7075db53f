"""Negative-gamma uniform case: error structure of the feature gradient for the three native kernel sets."""
import copy, os, sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_stack import pointnet2_modules
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
dev = torch.device("cuda:0")
for seed in (3, 4, 5, 6, 7, 8, 9, 10):
    torch.manual_seed(seed)
    Bp, S = 6, 512
    xyz = (torch.rand(Bp * S, 3, device=dev) - 0.5) * torch.tensor([0.8, 0.8, 1.8], device=dev)
    feats = torch.rand(Bp * S, 1, device=dev)
    cnt = torch.full((Bp,), S, dtype=torch.int32, device=dev)
    native = pointnet2_modules.StackSAModuleMSG(radii=[0.3], nsamples=[32], mlps=[[1, 64, 128]]).to(dev)
    with torch.no_grad():
        for m in native.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.weight.normal_(0.3, 1.0); m.bias.normal_(0, 0.3)
    w_out = torch.randn((Bp * S, 128), device=dev)
    res = {}
    for name in ("tc", "fast", "generic", "torch"):
        os.environ["MGAR_SA_TRAIN"] = "torch" if name == "torch" else "native"
        os.environ["MGAR_SA_TRAIN_KERNELS"] = name
        mod = copy.deepcopy(native).train()
        f = feats.clone().requires_grad_(True)
        _, out = mod(xyz, cnt, xyz.clone(), cnt, f)
        (out * w_out).sum().backward()
        res[name] = (out.detach(), f.grad.clone(), [p.grad.clone() for p in mod.parameters()])
    ro, rg, rp = res["torch"]
    for name in ("tc", "fast", "generic"):
        o, g, ps = res[name]
        d = (g - rg).abs().flatten(); sc = float(rg.abs().max())
        print("seed", seed, name, "out %.2e" % (float((o - ro).abs().max()) / float(ro.abs().max())),
              "gfeat max %.2e" % (float(d.max()) / sc), "rows>1e-4: %d" % int((d > 1e-4 * sc).sum()), "rows>1e-5: %d" % int((d > 1e-5 * sc).sum()),
              "params", ["%.1e" % (float((a - b).abs().max()) / float(b.abs().max())) for a, b in zip(ps, rp)])
