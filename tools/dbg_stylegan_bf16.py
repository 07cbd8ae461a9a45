"""Diagnostic: where does the bf16 StyleGAN2 gradient error come from?  Ours vs fp32 CPU oracle vs stock-torch
bf16 witness, with the W-folding / fused modulation glue switched off in turn."""
import argparse, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200")); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import b200conv3d as B
import b200conv3d.functional as BF
import ref_stylegan as R
from test_stylegan_net import _set_noise_strength

res, cb, wd, zd, n = 64, 2048, 64, 32, 4
p = argparse.Namespace(z_size=zd, filterG=cb, filterD=cb, device="cuda", stylegan2=True, stylegan=False)
torch.manual_seed(3)
oD, oG = R.Discriminator(cb, img_resolution=res), R.Generator(zd, cb, w_dim=wd, img_resolution=res)
D, G = B.stylegan.Discriminator(p, img_resolution=res), B.stylegan.Generator(p, w_dim=wd, img_resolution=res)
_set_noise_strength(oG)
G.load_state_dict(oG.state_dict()), D.load_state_dict(oD.state_dict())
G, D = G.cuda(), D.cuda()
z = torch.randn(n, zd, 1, 1, 1)
real = torch.rand(n, 1, res, res, res) * 2 - 1
KEYS = ["G.synthesis.b64.conv1.weight", "G.synthesis.b32.conv1.weight", "G.synthesis.b64.conv0.weight", "G.synthesis.b16.conv1.weight",
        "G.synthesis.b8.conv1.weight", "D.b64.conv0.weight", "D.b32.conv1.weight", "G.mapping.fc7.weight"]

def run(g, d, dev, amp):
    for q in list(g.parameters()) + list(d.parameters()):
        q.grad = None
    ctx = torch.autocast("cuda", dtype=torch.bfloat16) if amp else torch.autocast("cuda", enabled=False)
    with ctx:
        img, _ = g(z.to(dev), noise_mode="const")
        img.retain_grad()
        lf = d(img)
        torch.nn.functional.softplus(lf.float()).mean().backward()
    out = {"img": img.detach().float().cpu().clone(), "dimg": img.grad.float().cpu().clone(), "lf": lf.detach().float().cpu().clone()}
    for k, q in list(("G." + k, q) for k, q in g.named_parameters()) + list(("D." + k, q) for k, q in d.named_parameters()):
        if k in KEYS:
            out[k] = q.grad.detach().float().cpu().clone()
    return out

l2 = lambda a, b: ((a - b).norm() / b.norm().clamp_min(1e-30)).item()
ref = run(oG, oD, "cpu", False)
rows = {"ours": run(G, D, "cuda", True)}
ms = BF.mod_supported
BF.mod_supported = lambda c, dt: False
rows["ours_noglue"] = run(G, D, "cuda", True)
BF.mod_supported = ms
# only torgb through the elementwise formulation
import b200conv3d.stylegan_ops as SO
of = B.stylegan.OutBlock.forward
def torgb_plain(self, x, w, fused_modconv=False):
    styles = self.affine(w) * self.weight_gain
    x = SO.modulated_conv3d(x=x, weight=self.weight, styles=styles, demodulate=False, fused_modconv=False)
    return SO.bias_act(x, self.bias)
B.stylegan.OutBlock.forward = torgb_plain
rows["torgb_plain"] = run(G, D, "cuda", True)
B.stylegan.OutBlock.forward = of
# only the synthesis layers through the elementwise formulation
sc = SO.synthesis_conv
def synth_plain(x, weight, styles, noise_field, noise_strength, bias, up=1, padding=1, resample_filter=None, act="lrelu", demodulate=True):
    nz = None if noise_field is None else noise_field * noise_strength
    y = SO.modulated_conv3d(x=x, weight=weight, styles=styles, noise=nz, up=up, padding=padding, resample_filter=resample_filter,
                            demodulate=demodulate, flip_weight=(up == 1), fused_modconv=False)
    return SO.bias_act(y, bias, act=act)
SO.synthesis_conv = synth_plain
rows["synth_plain"] = run(G, D, "cuda", True)
SO.synthesis_conv = sc
rows["ours_fp32"] = run(G, D, "cuda", False)
rows["witness"] = run(oG.cuda(), oD.cuda(), "cuda", True)
print("%-34s" % "relative L2 vs fp32 oracle" + "".join("%14s" % k for k in rows))
for k in ["img", "lf", "dimg"] + KEYS:
    print("%-34s" % k + "".join("%14.4f" % l2(rows[r][k], ref[k]) for r in rows))
