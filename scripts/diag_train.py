"""Relative errors (to the largest entry) of every output of the native training path against torch autograd."""
import copy, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_stack import pointnet2_modules
import cases
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
def run(mod, args, feats, w_out):
    f = feats.clone().requires_grad_(True)
    _, out = mod(args[0], args[1], args[2], args[3], f)
    (out * w_out).sum().backward()
    return [("out", out.detach()), ("gfeat", f.grad)] + [(n, p.grad.clone()) for n, p in mod.named_parameters()] + \
           [(n, b.clone()) for n, b in mod.named_buffers() if b.dtype.is_floating_point]
for cin in (2, 1):
    xyz, cnt, new_xyz, ncnt, feats = cases.ragged_case()
    feats = feats[:, :cin]
    torch.manual_seed(11)
    native = pointnet2_modules.StackSAModuleMSG(radii=[0.6], nsamples=[32], mlps=[[cin, 64, 128]], pool_method="max_pool").to(dev)
    with torch.no_grad():
        for m in native.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.weight.normal_(1, 0.2); m.bias.normal_(0, 0.2)
    args = (T(xyz), T(cnt), T(new_xyz), T(ncnt)); f = T(feats)
    w_out = torch.randn((new_xyz.shape[0], 128), device=dev)
    res = {}
    for name in ("fast", "generic", "torch"):
        os.environ["MGAR_SA_TRAIN"] = "torch" if name == "torch" else "native"
        os.environ["MGAR_SA_TRAIN_KERNELS"] = name
        mod = copy.deepcopy(native).train()
        res[name] = run(mod, args, f, w_out)
    for (n, a), (_, b), (_, r) in zip(res["fast"], res["generic"], res["torch"]):
        sc = max(1e-6, float(r.abs().max()))
        print("C_in=%d %-34s fast %.2e generic %.2e" % (cin, n, float((a - r).abs().max()) / sc, float((b - r).abs().max()) / sc))
    if cin == 2:
        fa, to = dict(res["fast"]), dict(res["torch"])
        d = (fa["mlps.0.1.bias"] - to["mlps.0.1.bias"]).abs()
        print("beta1 err by channel top:", torch.topk(d, 6), "scale", float(to["mlps.0.1.bias"].abs().max()))
        d = (fa["gfeat"] - to["gfeat"]).abs()
        v, i = torch.topk(d.flatten(), 12)
        print("gfeat err top:", v.tolist(), (i // 2).tolist(), (i % 2).tolist())
        print("gfeat rows with err>1e-4*max:", int((d.max(1).values > 1e-4 * float(to["gfeat"].abs().max())).sum()), "of", d.shape[0])
        d = (fa["mlps.0.0.weight"] - to["mlps.0.0.weight"]).abs().reshape(64, 5)
        print("dW1 err by input column:", d.max(0).values.tolist())
        print("dW1 err by out channel top:", torch.topk(d.max(1).values, 6))
