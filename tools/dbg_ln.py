import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))
import torch, torch.nn as nn
import b200conv3d as B
torch.manual_seed(3)
ref = nn.Sequential(nn.LayerNorm([16, 4, 6, 8]), nn.LeakyReLU(0.2))
with torch.no_grad():
    ref[0].weight.normal_(1, 0.2); ref[0].bias.normal_(0, 0.2)
ours = nn.Sequential(B.nn.LayerNorm([16, 4, 6, 8]), B.nn.LeakyReLU(0.2))
ours.load_state_dict(ref.state_dict()); ours.cuda(); B.fuse(ours)
x = torch.randn(5, 16, 4, 6, 8) * 1.5 + 0.3
g = torch.randn(5, 16, 4, 6, 8)
for mode in ("fp32", "bf16", "bf16-inputs-prerounded"):
    xx = x.bfloat16().float() if "pre" in mode else x
    gg = g.bfloat16().float() if "pre" in mode else g
    xr = xx.clone().requires_grad_(True); xo = xx.clone().cuda().requires_grad_(True)
    yr = ref(xr)
    if mode == "fp32":
        yo = ours(xo)
    else:
        with torch.autocast("cuda", dtype=torch.bfloat16):
            yo = ours(xo)
    yr.backward(gg); yo.backward(gg.cuda().to(yo.dtype))
    e = (xo.grad.cpu().double() - xr.grad.double()).abs() / xr.grad.abs().max()
    ey = (yo.float().cpu().double() - yr.double()).abs() / yr.abs().max()
    print(mode, "y err max %.2e" % ey.max().item(), "dx err max %.2e mean %.2e frac>2e-2 %.4f" % (e.max().item(), e.mean().item(), (e > 2e-2).double().mean().item()),
          "max|dx| %.3f" % xr.grad.abs().max().item())
    for n1, (p1, p2) in zip(("gamma", "beta"), zip(ref.parameters(), ours.parameters())):
        ep = (p2.grad.cpu().double() - p1.grad.double()).abs() / p1.grad.abs().max()
        print("   ", n1, "err max %.2e mean %.2e" % (ep.max().item(), ep.mean().item()))
        p1.grad = None; p2.grad = None
    # where are the outliers
    idx = (e > 2e-2).nonzero()
    if len(idx):
        print("   first outliers", idx[:6].tolist(), "per-sample counts", [(e[i] > 2e-2).sum().item() for i in range(5)])
        i = tuple(idx[0].tolist())
        print("   ours %.4f ref %.4f" % (xo.grad.cpu()[i].item(), xr.grad[i].item()))
