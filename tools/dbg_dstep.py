import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import torch, ref_dcgan
import b200conv3d as B
width, z, b = 2, 8, 2
torch.manual_seed(1)
rg, rd = ref_dcgan.Generator(z, width), ref_dcgan.Discriminator(width)
g, d = B.dcgan.build(z, width, width, 128, "cpu", fuse=False)
g.load_state_dict(rg.state_dict()); d.load_state_dict(rd.state_dict())
g, d = g.cuda(), d.cuda(); B.fuse(g); B.fuse(d)
zin = torch.randn(b, z, 1, 1, 1); real = torch.rand(b, 1, 128, 128, 128) * 2 - 1
# reference grads of D_step loss
for p in rd.parameters(): p.requires_grad = True
fake_r = rg(zin).detach()
(rd(fake_r).mean() - rd(real).mean()).backward()
for p in d.parameters(): p.requires_grad = True
for p in g.parameters(): p.requires_grad = False
with torch.no_grad():
    fake = g(zin.cuda())
print("fake err", ((fake.cpu() - fake_r).abs().max() / fake_r.abs().max()).item())
(d(fake).mean() - d(real.cuda()).mean()).backward()
for (n, p), q in zip(rd.named_parameters(), d.parameters()):
    e = (q.grad.cpu() - p.grad).abs().max().item(); s = p.grad.abs().max().item()
    sign_mismatch = ((q.grad.cpu().sign() != p.grad.sign()) & (p.grad.abs() > 1e-7)).float().mean().item()
    print("%-16s max|ref| %.3e  maxerr %.3e  rel %.2e  sign-mismatch(|g|>1e-7) %.4f  frac|g|<1e-8 %.4f" % (n, s, e, e / max(s, 1e-30), sign_mismatch, (p.grad.abs() < 1e-8).float().mean().item()))
