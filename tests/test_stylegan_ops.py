"""StyleGAN2 convolution funnel (b200conv3d/stylegan_ops.py: conv3d_resample, modulated_conv3d,
upfirdn3d) against golden vectors of the UNMODIFIED reference functions
(tests/golden/stylegan_ops_ref.npz, oracle/gen_golden_stylegan.py).

CPU: (1) the pinned oracle restatement replays the golden file; (2) the product's HOST logic
(padding dispatch, filter flips, demodulation GEMM) is run with the Python face bound to the
C oracle library in a subprocess.  GPU: the same cases through the CUDA kernels, fp32 (exact
kernels, 1e-4) and bf16 (tcgen05 / FIR kernels, 2e-2)."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ref_stylegan_ops as O  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden", "stylegan_ops_ref.npz")
FIR_CASES = (("fir_img_up2", dict(up=2, padding=[2, 1, 2, 1, 2, 1], gain=8)),
             ("fir_down2", dict(down=2, padding=[1, 1, 1, 1, 1, 1])),
             ("fir_pad", dict(padding=[2, 1, 2, 1, 2, 1])),
             ("fir_crop", dict(padding=[-1, -1, -1, -1, -1, -1])))


def _grads(ops, case, dev, dtype=None, autocast=False):
    name, kind = case[0], case[1]
    x, w, s, nz, gy = O.case_inputs(*case)
    x, w, s, gy = (t.to(dev) for t in (x, w, s, gy))
    nz = nz.to(dev) if nz is not None else None
    x.requires_grad_(True), w.requires_grad_(True), s.requires_grad_(True)
    if autocast:
        with torch.autocast("cuda", dtype=torch.bfloat16):
            y = O.run_case(ops, case, x.bfloat16(), w, s, nz)
        y.backward(gy.to(y.dtype))
    else:
        y = O.run_case(ops, case, x, w, s, nz)
        y.backward(gy)
    return {"y": y.detach(), "dx": x.grad, "dw": w.grad, "ds": s.grad if kind == "mod" else None}


def _check(got, gold, name, tol):
    for what, v in got.items():
        if v is None:
            continue
        ref = torch.from_numpy(gold["%s.%s" % (name, what)])
        err = ((v.float().cpu() - ref).abs().max() / ref.abs().max()).item()
        assert err < tol, (name, what, err)


@pytest.mark.parametrize("case", O.CASES, ids=[c[0] for c in O.CASES])
def test_oracle_replays_reference_golden(case):
    gold = np.load(GOLD)
    _check(_grads(O, case, "cpu"), gold, case[0], 1e-5)


def test_product_host_logic_on_oracle_library(oracle_path):
    """Runs b200conv3d.stylegan_ops on CPU tensors with the C ABI bound to the oracle library."""
    code = r'''
import sys, numpy as np, torch
sys.path.insert(0, %r); sys.path.insert(0, %r); sys.path.insert(0, %r)
from b200conv3d import _lib
_lib.use_library(%r, host=True)
import b200conv3d.stylegan_ops as P
import ref_stylegan_ops as O
import test_stylegan_ops as T
gold = np.load(T.GOLD)
for case in O.CASES:
    T._check(T._grads(P, case, "cpu"), gold, case[0], 1e-4)
import b200conv3d.functional as BF
f = P.make_resample_filter()
assert torch.equal(f, O.make_filter())
for name, kw in T.FIR_CASES:
    y = BF.upfirdn3d(torch.from_numpy(gold[name + ".x"]), f, **kw)
    assert (y - torch.from_numpy(gold[name + ".y"])).abs().max() < 1e-5, name
print("HOST-OK")
''' % (os.path.join(ROOT, "3d-gan_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests"), oracle_path)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert "HOST-OK" in out.stdout, out.stdout[-2000:] + out.stderr[-4000:]


@pytest.mark.gpu
@pytest.mark.parametrize("case", O.CASES, ids=[c[0] for c in O.CASES])
def test_gpu_stylegan_ops_fp32(case):
    import b200conv3d.stylegan_ops as P
    _check(_grads(P, case, "cuda"), np.load(GOLD), case[0], 1e-4)


@pytest.mark.gpu
@pytest.mark.parametrize("case", O.CASES, ids=[c[0] for c in O.CASES])
def test_gpu_stylegan_ops_bf16(case):
    import b200conv3d.stylegan_ops as P
    got = _grads(P, case, "cuda", autocast=True)
    _check(got, np.load(GOLD), case[0], 3e-2)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_gpu_fir_vs_reference_golden(dtype):
    import b200conv3d.functional as BF
    import b200conv3d.stylegan_ops as P
    gold = np.load(GOLD)
    f = P.make_resample_filter()
    for name, kw in FIR_CASES:
        x = torch.from_numpy(gold[name + ".x"]).cuda().to(dtype).requires_grad_(True)
        y = BF.upfirdn3d(x, f, **kw)
        ref = torch.from_numpy(gold[name + ".y"])
        tol = 1e-5 if dtype == torch.float32 else 2e-2
        assert ((y.float().cpu() - ref).abs().max() / ref.abs().max()).item() < tol, name
        # adjoint identity <fir(x), g> == <x, fir^T(g)> ties the backward kernel to the forward one
        g = torch.randn_like(y)
        y.backward(g)
        lhs = (y.detach().float() * g.float()).sum().item()
        rhs = (x.detach().float() * x.grad.float()).sum().item()
        scale = (y.detach().float().norm() * g.float().norm()).item()       # Cauchy-Schwarz bound of either side
        assert abs(lhs - rhs) < (1e-5 if dtype == torch.float32 else 1e-2) * scale, (name, lhs, rhs, scale)


FIR_SHAPES = (
    # name, (n, c, d, h, w), kwargs of upfirdn3d  -- ragged extents: odd w, d not a multiple of the 16 / 32 chunk
    ("unit_up_conv", (2, 32, 33, 17, 19), dict(padding=[1, 1, 1, 1, 1, 1], gain=4)),        # after an up conv (:281)
    ("unit_pre_down", (1, 16, 70, 12, 9), dict(padding=[2, 2, 2, 2, 2, 2])),               # before a stride-2 conv (:258)
    # the shared-memory tiled march (16 x 16 tiles, 32-plane chunks): ragged tiles, three d chunks, two channel groups
    ("unit_tiled", (1, 64, 70, 37, 41), dict(padding=[2, 2, 2, 2, 2, 2])),
    ("unit_tiled_crop", (2, 32, 21, 24, 35), dict(padding=[-1, -1, -1, -1, -1, -1])),
    ("unit_crop", (2, 8, 21, 10, 11), dict(padding=[-1, -1, -1, -1, -1, -1])),
    ("unit_odd_channels", (1, 3, 9, 8, 7), dict(padding=[2, 1, 2, 1, 2, 1])),
    ("decimate", (2, 16, 18, 14, 10), dict(down=2, padding=[1, 1, 1, 1, 1, 1])),             # skip path (:246)
    ("decimate_odd", (1, 32, 19, 13, 11), dict(down=2, padding=[2, 1, 2, 1, 2, 1])),        # odd extents, even low pad
    ("decimate_nopad", (1, 8, 12, 10, 9), dict(down=2, padding=[0, 0, 0, 0, 0, 0])),
    ("zero_insert", (2, 8, 6, 5, 7), dict(up=2, padding=[2, 1, 2, 1, 2, 1], gain=8)),        # image up-sampling (:634)
    ("zero_insert_1ch", (2, 1, 8, 8, 8), dict(up=2, padding=[2, 1, 2, 1, 2, 1], gain=8)),
)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("case", FIR_SHAPES, ids=[c[0] for c in FIR_SHAPES])
def test_gpu_fir_vs_oracle_ragged(case, dtype):
    """Every FIR kernel variant (separable sliding window, w4, generic with strided taps, adjoints)
    against the pinned torch restatement of `_upfirdn3d_ref` on ragged shapes, forward and gradient.
    A second, non-separable filter sends the unit-rate cases through the non-separable kernel."""
    import b200conv3d.functional as BF
    name, shape, kw = case
    g = torch.Generator().manual_seed(sum(map(ord, name)))
    x = torch.randn(*shape, generator=g)
    f_sep = O.make_filter()
    f_full = f_sep + 0.05 * torch.rand(4, 4, 4, generator=g)
    for f in (f_sep, f_full):
        xr = x.clone().requires_grad_(True)
        p = kw["padding"]
        yr = O.upfirdn3d(xr, f, up=kw.get("up", 1), down=kw.get("down", 1), pad=(p[0], p[1]), gain=kw.get("gain", 1))
        gy = torch.randn(*yr.shape, generator=g)
        yr.backward(gy)
        xg = x.cuda().to(dtype).requires_grad_(True)
        y = BF.upfirdn3d(xg, f, **kw)
        y.backward(gy.cuda().to(dtype))
        tol = 1e-5 if dtype == torch.float32 else 2e-2
        assert y.shape == yr.shape
        assert ((y.float().cpu() - yr.detach()).abs().max() / yr.abs().max()).item() < tol, name
        assert ((xg.grad.float().cpu() - xr.grad).abs().max() / xr.grad.abs().max()).item() < tol, name


MOD_SHAPES = ((2, 8, 5, 6, 7), (3, 32, 9, 4, 11), (2, 512, 4, 4, 4), (1, 64, 17, 3, 5))


def check_mod_ops(dev, dtype):
    """b2c_modulate / b2c_mod_epilogue (forward and every gradient) against the elementwise formulation of
    stylegan.py:427-432, 522-524, 531 written with stock torch ops in fp32."""
    import b200conv3d.functional as BF
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    err = lambda a, b: ((a.float().cpu() - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()   # noqa: E731
    for shape in MOD_SHAPES:
        n, c = shape[:2]
        g = torch.Generator().manual_seed(sum(shape))
        x = torch.randn(*shape, generator=g)
        s = torch.randn(n, c, generator=g) * 0.5 + 1
        gy = torch.randn(*shape, generator=g)
        xr, sr = x.clone().requires_grad_(True), s.clone().requires_grad_(True)
        (xr * sr.reshape(n, c, 1, 1, 1)).backward(gy)
        xg, sg = x.clone().to(dev).to(dtype).requires_grad_(True), s.clone().to(dev).requires_grad_(True)
        y = BF.modulate(xg, sg)
        y.backward(gy.to(dev).to(dtype))
        assert err(y, (x * s.reshape(n, c, 1, 1, 1))) < tol
        assert err(xg.grad, xr.grad) < tol and err(sg.grad, sr.grad) < tol, shape
        for demod, fn, act in ((True, n, "lrelu"), (True, 1, "lrelu"), (False, 0, None), (True, 0, "tanh")):
            d = torch.rand(n, c, generator=g) + 0.5 if demod else None
            field = torch.randn(fn, 1, *shape[2:], generator=g) if fn else None
            st = torch.tensor(0.3)
            b = torch.randn(c, generator=g) * 0.1
            ref_in = [t.clone().requires_grad_(True) if t is not None else None for t in (x, d, st if fn else None, b)]
            z = ref_in[0] * (ref_in[1].reshape(n, c, 1, 1, 1) if demod else 1.0)
            if fn:
                z = z + field * ref_in[2]
            z = z + ref_in[3].reshape(1, c, 1, 1, 1)
            ref = torch.nn.functional.leaky_relu(z, 0.2) if act == "lrelu" else (torch.tanh(z) if act == "tanh" else z)
            ref.backward(gy)
            got_in = [t.detach().clone().to(dev).requires_grad_(True) if t is not None else None
                      for t in (x.to(dtype), d, st if fn else None, b)]
            out = BF.mod_epilogue(got_in[0], got_in[1], None if field is None else field.to(dev), got_in[2], got_in[3],
                                  act=act, slope=0.2)
            out.backward(gy.to(dev).to(dtype))
            assert err(out, ref.detach()) < tol, (shape, demod, fn, act)
            for a, r, what in zip(got_in, ref_in, ("dy", "dd", "dstrength", "dbias")):
                if r is not None:
                    # bf16: the kernels take act' from the rounded OUTPUT; elements within rounding of a kink may flip
                    if dtype == torch.bfloat16 and what == "dy":
                        from abi import kink_robust_err
                        assert kink_robust_err(a.grad, r.grad, tol) < tol, (shape, demod, fn, act, what)
                    elif dtype == torch.bfloat16:
                        # signed sums over all voxels (and, behind lrelu, with a few kink flips): their bf16 error
                        # scales with the summands, not with the sum
                        assert (a.grad.float().cpu() - r.grad).abs().max().item() < tol * gy.norm().item(), \
                            (shape, demod, fn, act, what)
                    else:
                        assert err(a.grad, r.grad) < tol, (shape, demod, fn, act, what, err(a.grad, r.grad))


def test_mod_ops_oracle_library_vs_torch(oracle_path):
    code = r'''
import sys, torch
sys.path.insert(0, %r); sys.path.insert(0, %r); sys.path.insert(0, %r)
from b200conv3d import _lib
_lib.use_library(%r, host=True)
import test_stylegan_ops as T
T.check_mod_ops("cpu", torch.float32)
print("HOST-OK")
''' % (os.path.join(ROOT, "3d-gan_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests"), oracle_path)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert "HOST-OK" in out.stdout, out.stdout[-2000:] + out.stderr[-4000:]


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_gpu_mod_ops_vs_torch(dtype):
    check_mod_ops("cuda", dtype)


ADAIN_SHAPES = ((2, 8, 3, 4, 5), (3, 32, 6, 2, 7), (2, 512, 4, 4, 4), (1, 16, 1, 5, 3))


def check_adain(dev, dtype):
    """b2c_instance_stats / b2c_adain_* (with and without the folded trilinear x2) against the reference's
    formulation (stylegan.py:931-959 + F.interpolate(scale_factor=2, mode='trilinear')) in stock torch fp32."""
    import b200conv3d.functional as BF
    tol = 2e-5 if dtype == torch.float32 else 3e-2
    err = lambda a, b: ((a.float().cpu() - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()   # noqa: E731
    for shape in ADAIN_SHAPES:
        for up in (False, True):
            n, c = shape[:2]
            g = torch.Generator().manual_seed(sum(shape) + up)
            x = (torch.randn(*shape, generator=g) * 1.5 + 0.7).to(dtype).float()     # the same rounded input for both
            sm, ss = torch.randn(n, c, generator=g), torch.rand(n, c, generator=g) + 0.5
            xr, mr, sr = (t.clone().requires_grad_(True) for t in (x, sm, ss))
            flat = xr.reshape(n, c, -1)
            std = (flat.var(dim=2) + 1e-5).sqrt()[:, :, None, None, None] if flat.shape[2] > 1 else None
            if std is None:
                continue
            mean = flat.mean(dim=2)[:, :, None, None, None]
            yr = (xr - mean) / std * sr[:, :, None, None, None] + mr[:, :, None, None, None]
            if up:
                yr = torch.nn.functional.interpolate(yr, scale_factor=2, mode="trilinear")
            gy = torch.randn(*yr.shape, generator=g)
            yr.backward(gy)
            xg, mg, sg = (t.clone().to(dev).requires_grad_(True) for t in (x.to(dtype), sm, ss))
            y = BF.adain(xg, mg, sg, 1e-5, up)
            y.backward(gy.to(dev).to(dtype))
            assert y.shape == yr.shape
            assert err(y, yr.detach()) < tol, (shape, up, err(y, yr.detach()))
            for a, r, what in ((xg, xr, "dx"), (mg, mr, "dmean"), (sg, sr, "dstd")):
                assert err(a.grad, r.grad) < tol, (shape, up, what, err(a.grad, r.grad))


def test_adain_oracle_library_vs_torch(oracle_path):
    code = r'''
import sys, torch
sys.path.insert(0, %r); sys.path.insert(0, %r); sys.path.insert(0, %r)
from b200conv3d import _lib
_lib.use_library(%r, host=True)
import test_stylegan_ops as T
T.check_adain("cpu", torch.float32)
print("HOST-OK")
''' % (os.path.join(ROOT, "3d-gan_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests"), oracle_path)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert "HOST-OK" in out.stdout, out.stdout[-2000:] + out.stderr[-4000:]


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_gpu_adain_vs_torch(dtype):
    check_adain("cuda", dtype)


@pytest.mark.gpu
@pytest.mark.parametrize("cin,cout,res", [(64, 32, 12), (32, 16, 9), (128, 64, 6)])
def test_gpu_phase_merged_up_conv_matches_scatter_form_and_oracle(cin, cout, res):
    """The narrow up-sampling layers (stylegan.py:262-282: conv_transpose3d k3 s2 + FIR) run as ONE stride-1
    `Cin -> 8*Cout` 2x2x2 conv + depth-to-space on the even grid (functional.conv_transpose3d_s2_even).  Result
    and both gradients against the library's sub-pixel scatter form on the same bf16 inputs, and against the
    fp32 formulation of oracle/ref_stylegan_ops.py."""
    sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))
    import b200conv3d.functional as BF
    import b200conv3d.stylegan_ops as P
    g = torch.Generator().manual_seed(cin + cout + res)
    x = torch.randn(2, cin, res, res, res, generator=g)
    w = torch.randn(cout, cin, 3, 3, 3, generator=g) / (cin * 27) ** 0.5
    gy = torch.randn(2, cout, 2 * res, 2 * res, 2 * res, generator=g)
    f = P.make_resample_filter()

    def run(merge):
        BF.PHASE_MERGE[0] = merge
        try:
            xs = x.cuda().requires_grad_(True)
            ws = w.cuda().requires_grad_(True)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                y = P.conv3d_resample(xs, ws, f=f.cuda(), up=2, padding=1, flip_weight=False)
            y.backward(gy.cuda().to(y.dtype))
            return y.float().cpu(), xs.grad.float().cpu(), ws.grad.float().cpu()
        finally:
            BF.PHASE_MERGE[0] = True

    assert BF.phase_merge_applies(x.cuda().bfloat16(), w.transpose(0, 1).cuda())
    got, ref = run(True), run(False)
    xr, wr = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    yo = O.conv3d_resample(xr, wr, f=f, up=2, padding=1, flip_weight=False)
    yo.backward(gy)
    gold = (yo.detach(), xr.grad, wr.grad)
    assert got[0].shape == ref[0].shape == gold[0].shape
    l2 = lambda a, b: ((a - b).norm() / b.norm().clamp_min(1e-30)).item()   # noqa: E731
    for a, b, o, name in zip(got, ref, gold, ("y", "dx", "dw")):
        # bf16 tolerance: both forms round the conv result to bf16 once before the FIR
        assert l2(a, o) < max(1.2e-2, 1.5 * l2(b, o)), (name, l2(a, o), l2(b, o))
        assert l2(a, b) < 1.5e-2, (name, l2(a, b))


@pytest.mark.gpu
@pytest.mark.parametrize("cin,cout,res", [(32, 64, 13), (64, 128, 12), (8, 16, 7)])
def test_gpu_phase_merged_dgrad_of_down_conv(cin, cout, res):
    """dgrad of the narrow stride-2 3^3 convs (stylegan.py:257-260) = the same merged conv + depth-to-space,
    written into the (odd or even) input grid: against the sub-pixel scatter form and fp32 torch."""
    sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))
    import b200conv3d.functional as BF
    g = torch.Generator().manual_seed(cin + cout + res)
    x = torch.randn(2, cin, res, res, res, generator=g)
    w = torch.randn(cout, cin, 3, 3, 3, generator=g) / (cin * 27) ** 0.5
    xr = x.clone().requires_grad_(True)
    yr = torch.nn.functional.conv3d(xr, w, stride=2)
    gy = torch.randn(*yr.shape, generator=g)
    yr.backward(gy)

    def run(merge):
        BF.PHASE_MERGE[0] = merge
        try:
            xs = x.cuda().bfloat16().requires_grad_(True)
            ws = w.cuda().requires_grad_(True)
            y = BF.conv3d(xs, ws, None, 2, 0)
            y.backward(gy.cuda().bfloat16())
            return xs.grad.float().cpu(), ws.grad.float().cpu()
        finally:
            BF.PHASE_MERGE[0] = True

    (dx1, dw1), (dx0, dw0) = run(True), run(False)
    l2 = lambda a, b: ((a - b).norm() / b.norm().clamp_min(1e-30)).item()   # noqa: E731
    assert dx1.shape == xr.grad.shape
    assert l2(dx1, xr.grad) < max(8e-3, 1.5 * l2(dx0, xr.grad)), (l2(dx1, xr.grad), l2(dx0, xr.grad))
    assert l2(dx1, dx0) < 8e-3 and l2(dw1, dw0) < 1e-5     # (8 channels: the catch-all direct wgrad accumulates with atomics)


@pytest.mark.gpu
@pytest.mark.parametrize("c,shape", [(32, (2, 5, 6, 7)), (16, (1, 9, 4, 3)), (96, (3, 4, 4, 5)), (64, (1, 16, 16, 16))])
def test_gpu_pointwise_one_channel_convs(c, shape):
    """fromrgb (1 -> C) and torgb (C -> 1) 1x1x1 convs (stylegan.py:542-546,758) take the pointwise kernels of
    csrc/thin.cu (outer product / row dot / weighted column sum): forward with the fused bias + activation and
    all three gradients against fp32 torch on the same bf16-rounded operands."""
    sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))
    import b200conv3d.functional as BF
    n, d, h, w = shape
    g = torch.Generator().manual_seed(c + sum(shape))
    bf = lambda t: t.bfloat16().float()   # noqa: E731
    l2 = lambda a, b: ((a.float().cpu() - b).norm() / b.norm().clamp_min(1e-30)).item()   # noqa: E731
    for cin, cout, act in ((1, c, "lrelu"), (c, 1, None)):
        x = bf(torch.randn(n, cin, d, h, w, generator=g))
        wt = bf(torch.randn(cout, cin, 1, 1, 1, generator=g) / cin ** 0.5)
        b = torch.randn(cout, generator=g) * 0.1
        gy = bf(torch.randn(n, cout, d, h, w, generator=g))
        xr, wr, br = (t.clone().requires_grad_(True) for t in (x, wt, b))
        yr = torch.nn.functional.conv3d(xr, wr, br)
        if act:
            yr = torch.nn.functional.leaky_relu(yr, 0.2)
        yr.backward(gy)
        xs = x.cuda().bfloat16().requires_grad_(True)
        ws, bs = wt.cuda().requires_grad_(True), b.cuda().requires_grad_(True)
        y = BF.conv3d(xs, ws, bs, 1, 0, act=act, slope=0.2)
        y.backward(gy.cuda().bfloat16())
        assert y.dtype == torch.bfloat16 and y.shape == yr.shape
        assert l2(y, yr.detach()) < 4e-3, (cin, cout, l2(y, yr.detach()))
        # lrelu': taken from the rounded output; elements within rounding of the kink may flip -> L2 criteria
        assert l2(xs.grad, xr.grad) < 8e-3, (cin, cout, "dx", l2(xs.grad, xr.grad))
        assert l2(ws.grad, wr.grad) < 8e-3, (cin, cout, "dw", l2(ws.grad, wr.grad))
        assert l2(bs.grad, br.grad) < 8e-3, (cin, cout, "db", l2(bs.grad, br.grad))
