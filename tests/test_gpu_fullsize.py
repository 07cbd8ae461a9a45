"""GPU parity at BASELINE sizes (config 2: DCGAN width 128, 128^3, batch 32, and batch 64 = the
concatenated D pass of GanStep; config 3: one BigGAN f=64 block at 128^3, batch 16; config 5: the
FID layer4 bottleneck at batch 32).  VALUES, not properties:

  (i)  the product library through the C ABI vs stock torch fp32 on the CPU (the reference's own
       arithmetic: oneDNN) on single samples of the batch - first, middle, last - and a 2-sample
       weight gradient;
  (ii) vs stock torch fp32 on the GPU (cuDNN / ATen with TF32 OFF) on the whole tensor.

Tolerances are north_star's: bf16 <= 2e-2 per layer, relative to max|ref| of the tensor; stacked
slices state theirs where they are used.  Set B2C_FULLSIZE_JSON=<path> to dump every measured error.
"""
import json
import os
import sys

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

import b200conv3d as B
from abi import Abi, make_desc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
BF = torch.bfloat16
TOL = 2e-2
_REPORT = {}


@pytest.fixture(scope="module", autouse=True)
def _exact_fp32_reference():
    """The GPU witness is fp32 with TF32 off; the report is written when the module finishes."""
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old
    path = os.environ.get("B2C_FULLSIZE_JSON")
    if path:
        with open(path, "w") as f:
            json.dump(_REPORT, f, indent=1, sort_keys=True)


def _note(name, **kv):
    _REPORT.setdefault(name, {}).update({k: float(v) for k, v in kv.items()})
    print(name, " ".join("%s=%.3e" % (k, v) for k, v in kv.items()))


def _err(ours, ref):
    """(max|a-b| / max|b|, ||a-b|| / ||b||), chunked over dim 0 so that 2^31-element tensors never need an fp32 copy.
    `ours` / `ref` are indexable by sample and give same-shaped tensors."""
    mx = mref = 0.0
    sq = sqr = 0.0
    for i in range(len(ours)):
        a, b = ours[i].float(), ref[i].float().to(ours[i].device)
        d = a - b
        mx = max(mx, d.abs().max().item())
        mref = max(mref, b.abs().max().item())
        sq += d.double().pow(2).sum().item()
        sqr += b.double().pow(2).sum().item()
    return mx / max(mref, 1e-30), (sq / max(sqr, 1e-300)) ** 0.5


class _Ndhwc:
    """Per-sample NDHWC view of a logical-NCDHW reference tensor."""

    def __init__(self, t):
        self.t = t

    def __len__(self):
        return self.t.shape[0]

    def __getitem__(self, i):
        return self.t[i].permute(1, 2, 3, 0)


# name, transposed, cin, cout, in-spatial, batches, act, bias            reference site
CONV_CASES = [
    ("D.L1", False, 1, 128, 128, (32, 64), 0, False),      # dcgan.py:161  (batch 64: 2^31 output elements)
    ("D.L2", False, 128, 256, 64, (32, 64), 0, False),     # dcgan.py:165  heaviest launch of the step
    ("D.L5", False, 1024, 2048, 8, (32, 64), 0, False),    # dcgan.py:177
    ("G.L2", True, 2048, 1024, 4, (32,), 0, True),         # dcgan.py:51   134 M-parameter filter
    ("G.L5", True, 256, 128, 32, (32,), 0, True),          # dcgan.py:63
    ("G.L6", True, 128, 1, 64, (32,), 3, True),            # dcgan.py:67-68 bias + tanh in the epilogue
]


def _torch_fwd(tr, x, w, b):
    return F.conv_transpose3d(x, w, b, 2, 1) if tr else F.conv3d(x, w, b, 2, 1)


@pytest.mark.parametrize("case", [(c, n) for c in CONV_CASES for n in c[5]], ids=lambda c: "%s-b%d" % (c[0][0], c[1]))
def test_conv_layer_values_at_headline_size(product_path, case):
    (name, tr, cin, cout, sp, _, act, has_bias), n = case
    gpu = Abi(product_path, DEV)
    torch.manual_seed(len(name) * 131 + n)
    osp = sp * 2 if tr else sp // 2
    desc = make_desc(tr, (n, cin, sp, sp, sp), (n, cout, osp, osp, osp), (4,) * 3, (2,) * 3, (1,) * 3, (1,) * 3, BF)
    x = torch.randn(n, sp, sp, sp, cin, device=DEV, dtype=BF)
    dy = torch.randn(n, osp, osp, osp, cout, device=DEV, dtype=BF)
    wshape = (cin, cout, 4, 4, 4) if tr else (cout, cin, 4, 4, 4)
    w = torch.randn(wshape, device=DEV) / (cin * (8 if tr else 64)) ** 0.5       # O(1) outputs: taps per output x Cin
    b = torch.randn(cout, device=DEV) * 0.1 if has_bias else None

    y = gpu.conv_fwd(desc, x, w, b, act, 0.0)
    dx = gpu.conv_dgrad(desc, dy, w)
    dw, db = gpu.conv_wgrad(desc, x, dy, wshape, want_bias=has_bias)
    assert gpu.last_algo() == 2, "tensor-core path was not taken"
    torch.cuda.synchronize()

    # (ii) whole tensor vs stock torch fp32 on the GPU (TF32 off), same bf16-rounded inputs
    xr = x.permute(0, 4, 1, 2, 3).float().requires_grad_(True)
    wr = w.clone().requires_grad_(True)
    z = _torch_fwd(tr, xr, wr, None)
    gx, gw = torch.autograd.grad(z, (xr, wr), dy.permute(0, 4, 1, 2, 3).float())
    with torch.no_grad():
        yr = z if b is None else z + b.view(1, -1, 1, 1, 1)
        yr = torch.tanh(yr) if act == 3 else yr
    e = {}
    e["fwd"], e["fwd_l2"] = _err(y, _Ndhwc(yr))
    e["dgrad"], e["dgrad_l2"] = _err(dx, _Ndhwc(gx))
    e["wgrad"], e["wgrad_l2"] = _err([dw], [gw])
    if has_bias:
        e["dbias"], _ = _err([db], [dy.float().sum(dim=(0, 1, 2, 3))])
    del z, yr, gx, xr

    # (i) single samples vs stock torch fp32 on the CPU (first / middle / last; the last one only at batch 64)
    wc = w.cpu()
    for i in ((0, n // 2, n - 1) if n == 32 else (n - 1,)):
        xi = x[i:i + 1].permute(0, 4, 1, 2, 3).float().cpu().requires_grad_(True)
        zi = _torch_fwd(tr, xi, wc, None)
        (gxi,) = torch.autograd.grad(zi, xi, dy[i:i + 1].permute(0, 4, 1, 2, 3).float().cpu())
        zi = zi.detach() if b is None else zi.detach() + b.cpu().view(1, -1, 1, 1, 1)
        zi = torch.tanh(zi) if act == 3 else zi
        e["cpu_fwd_s%d" % i], _ = _err(y[i:i + 1], _Ndhwc(zi))
        e["cpu_dgrad_s%d" % i], _ = _err(dx[i:i + 1], _Ndhwc(gxi))
    if n == 32:   # 2-sample weight gradient (its own split-K plan) vs the CPU
        d2 = make_desc(tr, (2, cin, sp, sp, sp), (2, cout, osp, osp, osp), (4,) * 3, (2,) * 3, (1,) * 3, (1,) * 3, BF)
        dw2, _ = gpu.conv_wgrad(d2, x[:2].contiguous(), dy[:2].contiguous(), wshape, want_bias=False)
        x2 = x[:2].permute(0, 4, 1, 2, 3).float().cpu()
        w2 = wc.clone().requires_grad_(True)
        (gw2,) = torch.autograd.grad(_torch_fwd(tr, x2, w2, None), w2, dy[:2].permute(0, 4, 1, 2, 3).float().cpu())
        e["cpu_wgrad_2samples"], _ = _err([dw2], [gw2])
    _note("conv %s b%d" % (name, n), **e)
    bad = {k: v for k, v in e.items() if not v < TOL}
    assert not bad, bad


# name, transposed, cin, cout, in-spatial, batch, kind (1 = per channel for BatchNorm3d, 2 = per sample for LayerNorm)
STAT_CASES = [
    ("G.L1", True, 512, 2048, 1, 32, 1),       # dcgan.py:47-48   point GEMM: columns are (tap, channel)
    ("G.L2", True, 2048, 1024, 4, 32, 1),      # dcgan.py:51-52
    ("G.L5", True, 256, 128, 32, 32, 1),       # dcgan.py:63-64   swapped epilogue, 8.4 M voxels per channel
    ("D.L1", False, 1, 128, 128, 64, 2),       # dcgan.py:161-162 one-channel gather kernel, 33.5 M values per sample
    ("D.L2", False, 128, 256, 64, 64, 2),      # dcgan.py:165-166 CTA pairs
    ("D.L5", False, 1024, 2048, 8, 64, 2),     # dcgan.py:177-178 two samples per 128-row tile
]


@pytest.mark.parametrize("case", STAT_CASES, ids=lambda c: c[0])
def test_epilogue_statistics_at_headline_size(product_path, case):
    """b2c_fusion.stat_*: the BatchNorm3d / LayerNorm statistics a conv accumulates in its GEMM epilogue equal the sums
    of the tensor it stored, at BASELINE sizes (fp64 reference over the stored bf16 values)."""
    import ctypes as C
    from b200conv3d._lib import Fusion
    name, tr, cin, cout, sp, n, kind = case
    gpu = Abi(product_path, DEV)
    L = gpu.lib
    torch.manual_seed(len(name) + n)
    k, s_, p_ = (4, 1, 0) if sp == 1 else (4, 2, 1)
    osp = (sp - 1) * s_ - 2 * p_ + k if tr else (sp + 2 * p_ - k) // s_ + 1
    desc = make_desc(tr, (n, cin, sp, sp, sp), (n, cout, osp, osp, osp), (k,) * 3, (s_,) * 3, (p_,) * 3, (1,) * 3, BF)
    x = torch.randn(n, sp, sp, sp, cin, device=DEV, dtype=BF)
    wshape = (cin, cout, 4, 4, 4) if tr else (cout, cin, 4, 4, 4)
    w = torch.randn(wshape, device=DEV) / (cin * (8 if tr and sp > 1 else 1 if tr else 64)) ** 0.5
    bias = torch.randn(cout, device=DEV) * 0.3 if tr else None
    wp = gpu.pack(desc, 0, w)
    y = torch.empty(n, osp, osp, osp, cout, dtype=BF, device=DEV)
    st = torch.full((2, cout if kind == 1 else n), float("nan"), device=DEV)
    core = (L.b2c_conv3d_workspace(C.byref(desc), 0) + 255) & ~255
    ws = torch.empty(max(core + L.b2c_conv3d_stats_workspace(C.byref(desc), kind), 16), dtype=torch.uint8, device=DEV)
    fu = Fusion(gpu._p(bias), 0, 0.0, kind, st[0].data_ptr(), st[1].data_ptr())
    rc = L.b2c_conv3d_fwd(C.byref(desc), gpu._p(x), gpu._p(wp), gpu._p(y), C.byref(fu), gpu._p(ws), ws.numel(), gpu._st())
    assert rc == 0, L.b2c_strerror(rc)
    torch.cuda.synchronize()
    rs = torch.zeros(st.shape[1], dtype=torch.float64, device=DEV)
    rq = torch.zeros_like(rs)
    for i in range(n):                      # fp64 over the stored values, one sample at a time
        yi = y[i].double()
        if kind == 1:
            rs += yi.sum(dim=(0, 1, 2)); rq += (yi * yi).sum(dim=(0, 1, 2))
        else:
            rs[i] = yi.sum(); rq[i] = (yi * yi).sum()
    count = y.numel() // st.shape[1]
    mean_o, mean_r = st[0].double() / count, rs / count
    var_o, var_r = st[1].double() / count - mean_o ** 2, rq / count - mean_r ** 2
    e = {"mean_err_over_std": ((mean_o - mean_r).abs() / var_r.sqrt()).max().item(),
         "var_rel_err": ((var_o - var_r).abs() / var_r).max().item()}
    _note("epilogue stats %s b%d" % (name, n), **e)
    assert e["mean_err_over_std"] < 1e-5 and e["var_rel_err"] < 2e-5, e


@pytest.mark.parametrize("n,c,s", [(32, 128, 64), (64, 128, 64), (32, 256, 32), (64, 2048, 4)])
def test_layernorm_lrelu_values_at_headline_size(n, c, s):
    """dcgan.py:162-179 LayerNorm([C,D,H,W]) + LeakyReLU(0.2), forward and backward (dx, full-shape dgamma / dbeta)
    vs ATen fp32 on the GPU, sample by sample."""
    torch.manual_seed(n + c)
    x = (torch.randn(n, s, s, s, c, device=DEV) * 1.5 + 0.3).to(BF).permute(0, 4, 1, 2, 3)
    g = torch.randn(n, s, s, s, c, device=DEV).to(BF).permute(0, 4, 1, 2, 3)
    gamma = (1 + 0.2 * torch.randn(c, s, s, s, device=DEV)).requires_grad_(True)
    beta = (0.2 * torch.randn(c, s, s, s, device=DEV)).requires_grad_(True)
    xo = x.detach().requires_grad_(True)
    with torch.autocast("cuda", dtype=BF):
        y = B.functional.layer_norm_act(xo, gamma, beta, 1e-5, "lrelu", 0.2)
    y.backward(g)
    dgo, dbo = gamma.grad.clone(), beta.grad.clone()
    yr, dxr, zr = [], [], []
    dgr, dbr = torch.zeros_like(gamma, dtype=torch.float64), torch.zeros_like(beta, dtype=torch.float64)
    gam, bet = gamma.detach().clone().requires_grad_(True), beta.detach().clone().requires_grad_(True)
    for i in range(n):
        xi = x[i:i + 1].float().requires_grad_(True)
        zi = F.layer_norm(xi, [c, s, s, s], gam, bet, 1e-5)
        yi = F.leaky_relu(zi, 0.2)
        gxi, gg, gb = torch.autograd.grad(yi, (xi, gam, bet), g[i:i + 1].float())
        yr.append(yi.detach()[0].to(BF)), dxr.append(gxi[0].to(BF)), zr.append(zi.detach()[0].to(BF))
        dgr += gg.double()
        dbr += gb.double()
    e = {}
    e["y"], e["y_l2"] = _err([y[i] for i in range(n)], yr)
    _, e["dx_l2"] = _err([xo.grad[i] for i in range(n)], dxr)
    _, e["dgamma_l2"] = _err([dgo], [dgr])
    _, e["dbeta_l2"] = _err([dbo], [dbr])
    # LeakyReLU kink: the two sides compute mean / rstd in different summation orders (1e-7 apart), so of the 1e9
    # pre-activations the few hundred that lie within 1e-6 of zero take the other slope.  Every element of dx that is
    # off by more than the tolerance must be such an element; dgamma / dbeta (sums over the batch) inherit them.
    scale = max(t.float().abs().max().item() for t in dxr)
    flips = bad_not_kink = 0
    for i in range(n):
        bad = (xo.grad[i].float() - dxr[i].float()).abs() > TOL * scale
        nb = int(bad.sum())
        if nb:
            flips += nb
            bad_not_kink += int((zr[i][bad].abs() > 1e-4).sum())
    e["kink_flips"] = flips
    e["outliers_not_at_a_kink"] = bad_not_kink
    e["dgamma_outlier_frac"] = ((dgo.double() - dgr).abs() > TOL * dgr.abs().max()).double().mean().item()
    e["dbeta_outlier_frac"] = ((dbo.double() - dbr).abs() > TOL * dbr.abs().max()).double().mean().item()
    _note("layernorm b%d c%d s%d" % (n, c, s), **e)
    # y / dx references are rounded to bf16 as well (one ulp = 4e-3 of the value): the tolerance is on max|ref|
    assert e["y"] < TOL and e["dx_l2"] < TOL and e["dgamma_l2"] < TOL and e["dbeta_l2"] < TOL, e
    assert bad_not_kink == 0 and flips <= 1e-5 * x.numel(), e
    assert e["dgamma_outlier_frac"] <= 1e-4 and e["dbeta_outlier_frac"] <= 1e-4, e


@pytest.mark.parametrize("n,c,s", [(32, 128, 64), (32, 256, 32)])
def test_batchnorm_relu_values_at_headline_size(n, c, s):
    """dcgan.py:48-65 BatchNorm3d + ReLU in training mode (statistics over 8.4 M rows), fwd + bwd + running stats."""
    torch.manual_seed(n + c + 1)
    x = (torch.randn(n, s, s, s, c, device=DEV) * 1.5 + 0.3).to(BF).permute(0, 4, 1, 2, 3)
    g = torch.randn(n, s, s, s, c, device=DEV).to(BF).permute(0, 4, 1, 2, 3)
    gamma = (1 + 0.2 * torch.randn(c, device=DEV)).requires_grad_(True)
    beta = (0.2 * torch.randn(c, device=DEV)).requires_grad_(True)
    rm, rv = torch.zeros(c, device=DEV), torch.ones(c, device=DEV)
    xo = x.detach().requires_grad_(True)
    with torch.autocast("cuda", dtype=BF):
        y = B.functional.batch_norm_act(xo, gamma, beta, rm, rv, True, 0.1, 1e-5, "relu")
    y.backward(g)
    rm2, rv2 = torch.zeros(c, device=DEV), torch.ones(c, device=DEV)
    gam, bet = gamma.detach().clone().requires_grad_(True), beta.detach().clone().requires_grad_(True)
    xr = x.float().requires_grad_(True)
    yr = F.relu(F.batch_norm(xr, rm2, rv2, gam, bet, True, 0.1, 1e-5))
    gx, gg, gb = torch.autograd.grad(yr, (xr, gam, bet), g.float())
    e = {}
    e["y"], e["y_l2"] = _err(y, yr.detach())
    e["dx"], e["dx_l2"] = _err(xo.grad, gx)
    e["dgamma"], _ = _err([gamma.grad], [gg])
    e["dbeta"], _ = _err([beta.grad], [gb])
    e["running_mean"], _ = _err([rm], [rm2])
    e["running_var"], _ = _err([rv], [rv2])
    _note("batchnorm b%d c%d s%d" % (n, c, s), **e)
    assert all(v < TOL for v in e.values()), e
    assert e["running_mean"] < 1e-3 and e["running_var"] < 1e-3


WIT = 1.25     # stacked bf16 slices: no worse than 1.25x what stock torch layers under bf16 autocast (cuDNN) lose


def _grad_report(tag, ours, ref, wit=None):
    """Relative L2 error of every parameter gradient of `ours` (and of the witness `wit`: the same stock layers under
    bf16 autocast) against the fp32 stock model `ref`.  A gradient that is mathematically zero - the bias of a conv
    that feeds a BatchNorm: fp32 gives 1e-7 of the largest gradient, any bf16 run the rounding noise of a 3e7-term
    sum - has no relative error; there the norm itself is compared with the witness's."""
    gmax = max(p.grad.abs().max().item() for p in ref.parameters())
    rows = {}
    wparams = list(wit.parameters()) if wit is not None else [None] * len(list(ref.parameters()))
    for (k, po), pr, pw in zip(ours.named_parameters(), ref.parameters(), wparams):
        assert po.grad is not None, k
        a, b = po.grad.double(), pr.grad.double().to(po.grad.device)
        w = pw.grad.double() if pw is not None else None
        if b.norm().item() < 1e-4 * gmax * b.numel() ** 0.5:
            rows[k] = ("zero", a.norm().item(), w.norm().item() if w is not None else float("nan"))
        else:
            # fewer than 64 elements (dcgan.py:67: the one bias of ConvT(128 -> 1)): the "norm" is a handful of draws
            rows[k] = ("l2" if b.numel() >= 64 else "few", ((a - b).norm() / b.norm()).item(),
                       ((w - b).norm() / b.norm()).item() if w is not None else float("nan"))
    _note(tag, **{k: v[1] for k, v in rows.items()})
    if wit is not None:
        _note(tag + " (witness)", **{k: v[2] for k, v in rows.items()})
    return rows


def _assert_grads(rows):
    def ok(v):
        if v[0] == "l2":
            return v[1] <= max(TOL, WIT * v[2])
        if v[0] == "few":                      # a scalar's error is one sample of the rounding noise, not a statistic
            return v[1] <= max(2.5 * TOL, 3 * v[2])
        return v[1] <= 2 * v[2] + 1e-30
    bad = {k: v for k, v in rows.items() if not ok(v)}
    assert not bad, bad


def test_dcgan_train_step_values_full_width(product_path):
    """trainer.py:208-283 at BASELINE config 2 (width 128, z 512, batch 32): one D_step + one G_step of the product
    path (bf16 autocast, the 64-sample concatenated D pass included) vs the same step of stock torch layers in fp32
    on the GPU (oracle/ref_dcgan.py = dcgan.py:45-70,158-182 restated, pinned to the reference by golden vectors),
    same weights, noise and real batch.  Per-sample logits, D losses and the generated volumes: <= 2e-2 (max, relative
    to max|logit|; the volumes in relative L2); the G loss, which crosses both networks, <= 4e-2.  Parameter gradients (6 to 24 stacked bf16 layers): relative L2 <= 2e-2,
    or <= 1.25x the error of the witness - the same stock layers under torch's bf16 autocast on this GPU."""
    import copy
    import ref_dcgan
    z, width, b = 512, 128, 32
    torch.manual_seed(1)
    rg, rd = ref_dcgan.Generator(z, width).to(DEV), ref_dcgan.Discriminator(width).to(DEV)
    with torch.no_grad():          # non-trivial affine parameters (init is gamma 1 / beta 0)
        for m in rd.modules():
            if isinstance(m, nn.LayerNorm):
                m.weight.normal_(1, 0.1), m.bias.normal_(0, 0.1)
    wg, wd = copy.deepcopy(rg), copy.deepcopy(rd)
    g, d = B.dcgan.build(z, width, width, 128, DEV, fuse=False)
    g.load_state_dict(rg.state_dict()), d.load_state_dict(rd.state_dict())
    B.fuse(g), B.fuse(d)
    real = torch.rand(b, 1, 128, 128, 128, device=DEV) * 2 - 1
    zd, zg = torch.randn(b, z, 1, 1, 1, device=DEV), torch.randn(b, z, 1, 1, 1, device=DEV)

    with torch.no_grad():
        lr_ref = rd(real).reshape(-1)
        fr = rg(zd)
        with torch.autocast("cuda", dtype=BF):
            lo = d(real).reshape(-1).float()
            fo = g(zd)
            lw = wd(real).reshape(-1).float()
            fw = wg(zd)
    scale = lr_ref.abs().max().item()
    e = {"logits_real": ((lo - lr_ref).abs().max() / scale).item()}
    ew = {"logits_real": ((lw - lr_ref).abs().max() / scale).item()}
    _, e["fake_l2"] = _err(fo, fr)
    _, ew["fake_l2"] = _err(fw, fr)
    del fo, fr, fw

    ref, wit = ref_dcgan.RefTrainer(rg, rd, z), ref_dcgan.RefTrainer(wg, wd, z)
    step = B.train.GanStep(g, d, z)
    assert step.batch_real_fake
    noise = [zd, zg]
    step.noise = lambda n_: noise.pop(0)
    r_real, r_fake = ref.d_step(real, zd)
    with torch.autocast("cuda", dtype=BF):
        w_real, w_fake = wit.d_step(real, zd)
    o_real, o_fake = step.D_step(real)
    e["errD_real"], e["errD_fake"] = abs(o_real.item() - r_real) / scale, abs(o_fake.item() - r_fake) / scale
    ew["errD_real"], ew["errD_fake"] = abs(w_real - r_real) / scale, abs(w_fake - r_fake) / scale
    gd = _grad_report("dcgan D_step grads", d, rd, wd)
    # G_step on identical discriminators: Adam's first step is sign-like, so the D's differ after it
    d.load_state_dict(rd.state_dict()), wd.load_state_dict(rd.state_dict())
    B.functional.invalidate_packs()
    r_g, _ = ref.g_step(b, zg)
    with torch.autocast("cuda", dtype=BF):
        w_g, _ = wit.g_step(b, zg)
    o_g, _ = step.G_step(b)
    e["errG"], ew["errG"] = abs(o_g.item() - r_g) / scale, abs(w_g - r_g) / scale
    gg = _grad_report("dcgan G_step grads", g, rg, wg)
    _note("dcgan step", **e)
    _note("dcgan step (witness)", **ew)
    # 2e-2 per network traversed: the D losses and logits pass through D (6 layers), errG through G and D (12 layers)
    assert all(v < (2 * TOL if k == "errG" else TOL) for k, v in e.items()), e
    _assert_grads(gd), _assert_grads(gg)


def _block_check(tag, ours, ref, x):
    """Forward: relative L2 <= 2e-2 and max <= 4e-2 of max|ref|.  Input / parameter gradients for a RANDOM upstream
    gradient (the worst case for a relative error: every per-channel sum is a random walk, and each ReLU whose
    pre-activation flips sign under bf16 rounding adds a full-size term): relative L2 <= 2e-2 or <= 1.25x the witness."""
    import copy
    ours.load_state_dict(ref.state_dict())
    wit = copy.deepcopy(ref)
    xr = x.float().requires_grad_(True)
    yr = ref(xr)
    gy = torch.randn_like(yr)
    yr.backward(gy)
    xw = x.detach().requires_grad_(True)
    with torch.autocast("cuda", dtype=BF):
        yw = wit(xw)
    yw.backward(gy.to(yw.dtype))
    xo = x.detach().requires_grad_(True)
    with torch.autocast("cuda", dtype=BF):
        yo = ours(xo)
    yo.backward(gy.to(yo.dtype))
    e, ew = {}, {}
    e["y"], e["y_l2"] = _err(yo, yr.detach())
    ew["y"], ew["y_l2"] = _err(yw, yr.detach())
    _, e["dx_l2"] = _err(xo.grad, xr.grad)
    _, ew["dx_l2"] = _err(xw.grad, xr.grad)
    _note(tag, **e)
    _note(tag + " (witness)", **ew)
    gr = _grad_report(tag + " grads", ours, ref, wit)
    assert e["y_l2"] < TOL and e["y"] < 2 * TOL, e
    assert e["dx_l2"] <= max(TOL, WIT * ew["dx_l2"]), (e, ew)
    _assert_grads(gr)


def test_biggan_blocks_values_at_128cubed():
    """BASELINE config 3 (run.sh:1, f = 64, batch 16): the last generator block (utils.py:47-86, 128 -> 64 channels,
    64^3 -> 128^3, hidden 32 channels at 128^3) and the first discriminator block (utils.py:88-132, 64 -> 128, hidden
    32 at 128^3, AvgPool) vs the stock-torch restatement in fp32 on the GPU; spectral norm on both sides."""
    import functools
    import ref_biggan
    torch.manual_seed(2)
    up = functools.partial(F.interpolate, scale_factor=2)
    rg = ref_biggan.GBlockDeep(128, 64, upsample=up).to(DEV)
    og = B.biggan.GBlockDeep(128, 64, upsample=up).to(DEV)
    x = torch.randn(16, 64, 64, 64, 128, device=DEV).to(BF).permute(0, 4, 1, 2, 3)
    _block_check("biggan GBlockDeep 128->64 @128^3", og, rg, x)
    del rg, og, x
    torch.cuda.empty_cache()
    rd = ref_biggan.DBlockDeep(64, 128, downsample=nn.AvgPool3d(2)).to(DEV)
    od = B.biggan.DBlockDeep(64, 128, downsample=B.nn.AvgPool3d(2)).to(DEV)
    x = torch.randn(16, 128, 128, 128, 64, device=DEV).to(BF).permute(0, 4, 1, 2, 3)
    _block_check("biggan DBlockDeep 64->128 @128^3", od, rd, x)


def test_fid_layer4_bottleneck_values_at_batch32():
    """BASELINE config 5 (FID_ResNet.py:65-102,136): a layer4 bottleneck, 2048 -> 512 -> (3^3, dilation 4) -> 2048 at
    16^3, batch 32, train-mode BatchNorm, forward + backward vs the fp32 restatement on the GPU."""
    import ref_fid_resnet
    torch.manual_seed(3)
    rb = ref_fid_resnet.Bottleneck(2048, 512, dilation=4).to(DEV)
    ob = B.fid_resnet.Bottleneck(2048, 512, dilation=4).to(DEV)
    x = torch.randn(32, 16, 16, 16, 2048, device=DEV).to(BF).permute(0, 4, 1, 2, 3)
    _block_check("fid layer4 bottleneck b32", ob, rb, x)
