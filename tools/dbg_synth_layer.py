"""Diagnostic: one SynthesisLayer body in bf16, fused glue vs elementwise formulation vs fp32."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))
import b200conv3d.functional as BF
import b200conv3d.stylegan_ops as SO
torch.manual_seed(0)
n, c, r = 4, 64, 32
x = torch.randn(n, c, r, r, r, device="cuda")
w = torch.randn(c, c, 3, 3, 3, device="cuda")
s = torch.randn(n, c, device="cuda") * 0.3 + 1
field = torch.randn(n, 1, r, r, r, device="cuda")
st = torch.tensor(0.1, device="cuda")
b = torch.randn(c, device="cuda") * 0.1
gy = torch.randn(n, c, r, r, r, device="cuda")
l2 = lambda a, b_: ((a.float() - b_.float()).norm() / b_.float().norm()).item()

def run(amp, glue):
    ms = BF.mod_supported
    if not glue:
        BF.mod_supported = lambda cc, dt: False
    ins = [t.clone().requires_grad_(True) for t in (x, w, s, st, b)]
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
        xin = ins[0].bfloat16() if amp else ins[0]
        y = SO.synthesis_conv(xin, ins[1], ins[2], field, ins[3], ins[4], up=1, padding=1, resample_filter=None, act="lrelu")
    y.backward(gy.to(y.dtype))
    BF.mod_supported = ms
    return [y.detach()] + [t.grad for t in ins]

ref = run(False, True)
names = ["y", "dx", "dw", "ds", "dstrength", "dbias"]
for tag, res in (("fp32 noglue", run(False, False)), ("bf16 fused", run(True, True)), ("bf16 noglue", run(True, False))):
    print("%-12s" % tag + "  ".join("%s %.4f" % (k, l2(a, r_)) for k, a, r_ in zip(names, res, ref)))
# decompose dw: wgrad term vs demod term
def terms(amp, glue):
    ms = BF.mod_supported
    if not glue:
        BF.mod_supported = lambda cc, dt: False
    w1 = w.clone().requires_grad_(True)
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
        xin = x.bfloat16() if amp else x
        d = SO.demod_coefficients(w1, s)
        y = SO.synthesis_conv(xin, w1, s, field, st, b, up=1, padding=1, resample_filter=None, act="lrelu")
    # total
    y.backward(gy.to(y.dtype))
    BF.mod_supported = ms
    return w1.grad
print("dw fused vs noglue (bf16): %.4f" % l2(terms(True, True), terms(True, False)))
