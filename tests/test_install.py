"""Face 1 of the boundary: the UNMODIFIED reference files build and run on the B200 layers after
``b200conv3d.install()``, ``convert()`` / ``prepare()`` retrofit stock-built models, state_dicts interchange, and the
reference's own ``Trainer.D_step`` / ``G_step`` (trainer.py:208-283, its fp16 ``autocast`` + ``GradScaler`` included)
reproduce the golden vectors.

* CPU (this container, /root/reference present): every check runs in a subprocess with the package bound to the
  plain-C oracle library (host pointers) - the layer plumbing under test is the product's, the arithmetic the oracle's.
* GPU (``-m gpu``): the same checks on cuda:0 through libb2c.so.  /root/reference does not exist on the GPU box, so
  there the model files are the pinned stock-torch restatements oracle/ref_*.py (they only say ``nn.Conv3d(...)``,
  like the reference) and the reference-tree checks are skipped with a reason.
"""
import argparse
import functools
import os
import subprocess
import sys
import tempfile
import types

import numpy as np
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = next((p for p in ("/root/reference", os.path.join(ROOT, "baseline", "_ref"))
            if os.path.exists(os.path.join(p, "dcgan.py"))), None)
needs_ref = pytest.mark.skipif(REF is None, reason="the reference tree (/root/reference) is not on this machine")


# ------------------------------------------------------------------------------------------------ helpers
def stub_modules():
    """The three arithmetic-free modules the reference imports but does not ship / need."""
    msl = types.ModuleType("msl")
    msl.RandomCrop3D = object
    fid = types.ModuleType("pytorch_fid_wrapper")
    fid.set_config = lambda **k: None
    fid.fid = lambda *a, **k: 0.0
    ct, ctt = types.ModuleType("carbontracker"), types.ModuleType("carbontracker.tracker")
    ctt.CarbonTracker = object
    ct.tracker = ctt
    sys.modules.update({"msl": msl, "pytorch_fid_wrapper": fid, "carbontracker": ct, "carbontracker.tracker": ctt})


def ref_import(name):
    stub_modules()
    sys.dont_write_bytecode = True
    if REF not in sys.path:
        sys.path.insert(0, REF)
    return __import__(name)


class installed:
    def __enter__(self):
        import b200conv3d as B
        B.install()

    def __exit__(self, *a):
        import b200conv3d as B
        B.uninstall()


def _rel(a, b, floor=0.0):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return ((a - b).abs().max() / max(b.abs().max().item(), floor, 1e-30)).item()


def _layers_are_ours(model):
    import b200conv3d as B
    pairs = ((nn.Conv3d, B.nn.Conv3d), (nn.ConvTranspose3d, B.nn.ConvTranspose3d), (nn.BatchNorm3d, B.nn.BatchNorm3d),
             (nn.LayerNorm, B.nn.LayerNorm), (nn.MaxPool3d, B.nn.MaxPool3d), (nn.AvgPool3d, B.nn.AvgPool3d))
    n = 0
    for m in model.modules():
        for stock, ours in pairs:
            if isinstance(m, stock):
                assert isinstance(m, ours), type(m)
                n += 1
    assert n > 0
    return n


def same_as_stock(make, inputs, dev, tol, train_inputs=True, fwd=None, functional_only=False):
    """`make()` builds a model from a reference file.  Built under install() it must consist of B200 layers, take the
    stock model's state_dict, and give the stock model's outputs, input gradients and parameter gradients."""
    fwd = fwd or (lambda m, *x: m(*x))
    torch.manual_seed(0)
    ref = make()                                     # stock torch layers
    with installed():
        torch.manual_seed(0)
        ours = make()
        if not functional_only:
            _layers_are_ours(ours)
        assert list(ours.state_dict()) == list(ref.state_dict())
        for (k, a), b in zip(ours.state_dict().items(), ref.state_dict().values()):
            assert a.shape == b.shape and a.dtype == b.dtype, k
            assert torch.equal(a, b), k               # same RNG draws: construction is arithmetic-identical
        ours = ours.to(dev)
        xs = [x.clone().to(dev).requires_grad_(train_inputs and x.is_floating_point()) for x in inputs]
        import unittest.mock
        import b200conv3d.functional as BF
        with unittest.mock.patch.object(BF, "conv3d", wraps=BF.conv3d) as spy:
            yo = fwd(ours, *xs)
        assert spy.call_count > 0 or not functional_only       # F.conv3d of the reference file reached the library
        g = torch.randn(yo.shape, generator=torch.Generator().manual_seed(1))
        yo.backward(g.to(dev))
    xr = [x.clone().requires_grad_(train_inputs and x.is_floating_point()) for x in inputs]
    yr = fwd(ref, *xr)
    yr.backward(g)
    exact = {}

    def fp64():
        """The same stock layers in double precision - the arbiter when stock fp32 itself is inexact: ATen's fp32
        LayerNorm backward over the 0.5-1 M-element rows of dcgan.py:162-166 is 3e-3..8e-3 away from it."""
        if not exact:
            import copy
            r64 = copy.deepcopy(ref).double()
            for q in r64.parameters():
                q.grad = None
            x64 = [x.clone().double().requires_grad_(train_inputs) if x.is_floating_point() else x.clone() for x in inputs]
            fwd(r64, *x64).backward(g.double())
            exact["x"], exact["p"] = [x.grad for x in x64], [q.grad for q in r64.parameters()]
        return exact

    def close(a, b, b64, tol_, floor=0.0):
        """within tol of stock fp32; or within tol of the exact (fp64) value; or - deep stacks with 2-sample batch
        statistics amplify fp32 rounding on both sides - no further from the exact value than 2x stock fp32 is."""
        if _rel(a, b, floor) < tol_:
            return True
        e64 = _rel(a, b64(), floor)
        return e64 < tol_ or e64 <= 2 * _rel(b, b64(), floor)

    assert _rel(yo, yr) < tol, ("forward", _rel(yo, yr))
    for i, (a, b) in enumerate(zip(xs, xr)):
        if b.grad is not None:
            assert close(a.grad, b.grad, lambda: fp64()["x"][i], tol), ("input gradient", _rel(a.grad, b.grad))
    gmax = max(p.grad.abs().max().item() for p in ref.parameters() if p.grad is not None)
    for i, ((k, po), pr) in enumerate(zip(ours.named_parameters(), ref.parameters())):
        if pr.grad is None:
            assert po.grad is None, k
            continue
        assert po.grad is not None, k
        assert close(po.grad, pr.grad, lambda: fp64()["p"][i], 10 * tol, 1e-2 * gmax), (k, _rel(po.grad, pr.grad, 1e-2 * gmax))
    # state_dict round trip back into the stock model (trainer.py:141-142,156-157; eval.py:26-27)
    ref.load_state_dict({k: v.cpu() for k, v in ours.state_dict().items()})
    return ours, ref


def _ns(**kw):
    d = dict(z_size=8, filterG=2, filterD=2, ngpu=1, sagan=False, sngan=False, msl=False, biggan=False, hinge=False,
             device="cpu", stylegan2=False, stylegan=False, dcgan=True, hybrid=False)
    d.update(kw)
    return argparse.Namespace(**d)


# ------------------------------------------------------------------------------------------------ checks
def check_reference_dcgan(dev, tol):
    """dcgan.py:8-195 unmodified: default G / LayerNorm D, and the --sngan D (row a3: SN + LeakyReLU(0.1), no LN)."""
    M = ref_import("dcgan")
    z = torch.randn(2, 8, 1, 1, 1, generator=torch.Generator().manual_seed(2))
    g, _ = same_as_stock(lambda: M.Generator(_ns()), [z], dev, tol)
    with torch.no_grad(), installed():
        img = g(z.to(dev)).float().cpu()
    assert tuple(img.shape) == (2, 1, 128, 128, 128)
    # D on a uniform volume in the data's clip range (data_handler.py:14): on G's near-constant output at init the
    # stock fp32 layers themselves are 8e-2 away from their fp64 result (a LeakyReLU kink in the 4^3 layer)
    img = torch.rand(2, 1, 128, 128, 128, generator=torch.Generator().manual_seed(9)) * 2 - 1
    same_as_stock(lambda: M.Discriminator(_ns()), [img], dev, tol)
    same_as_stock(lambda: M.Discriminator(_ns(sngan=True)), [img], dev, tol)


def check_reference_biggan_blocks(dev, tol):
    """utils.py:47-132 + biggan.py:8-119 unmodified: F.interpolate / nn.AvgPool3d are rebound too (no NCDHW bounce)."""
    U = ref_import("utils")
    Bg = ref_import("biggan")
    up = lambda: functools.partial(F.interpolate, scale_factor=2)       # evaluated at build time, like biggan.py:26
    gen = torch.Generator().manual_seed(3)
    same_as_stock(lambda: U.GBlockDeep(16, 8, upsample=up()), [torch.randn(2, 16, 4, 4, 4, generator=gen)], dev, tol)
    same_as_stock(lambda: U.DBlockDeep(8, 16, downsample=nn.AvgPool3d(2)), [torch.randn(2, 8, 8, 8, 8, generator=gen)],
                  dev, tol)
    # the default ResBlock generator (one block per stage, trainer.py:67-68) end to end at width 2, batch 2
    z = torch.randn(2, 8, generator=gen)
    same_as_stock(lambda: Bg.Generator(_ns(dcgan=False)), [z], dev, 5 * tol, train_inputs=False)


def check_reference_fid_resnet(dev, tol):
    """FID_ResNet.py unmodified: resnet50() forward as eval_utils.fid_3d runs it (no_grad, BatchNorm in train mode)."""
    M = ref_import("FID_ResNet")
    torch.manual_seed(4)
    ref = M.resnet50()
    with installed():
        torch.manual_seed(4)
        ours = M.resnet50()
        _layers_are_ours(ours)
        ours.load_state_dict(ref.state_dict())
        ours = ours.to(dev)
        x = torch.randn(2, 1, 32, 32, 32, generator=torch.Generator().manual_seed(5))
        with torch.no_grad():
            fo = ours(x.to(dev)).float().mean(dim=(2, 3, 4)).cpu()
    with torch.no_grad():
        fr = ref(x).mean(dim=(2, 3, 4))
    assert tuple(fo.shape) == (2, 2048)
    assert _rel(fo, fr) < 20 * tol, _rel(fo, fr)            # 53 stacked layers with 2-sample batch statistics


def check_reference_stylegan(dev, tol):
    """stylegan.py unmodified under install(): every F.conv3d / F.conv_transpose3d of `_conv3d_wrapper` (:164-170) and
    the depthwise `F.conv3d(x, f, groups=C)` FIR of `_upfirdn3d_ref` (:197) run in the library - no rebind needed."""
    S = ref_import("stylegan")
    p = argparse.Namespace(z_size=16, filterG=128, filterD=128, device="cpu", stylegan2=True, stylegan=False)
    z = torch.randn(2, 16, 1, 1, 1, generator=torch.Generator().manual_seed(6))
    g, _ = same_as_stock(lambda: S.Generator(p, w_dim=32, img_resolution=16), [z], dev, tol, train_inputs=False,
                         fwd=lambda m, zz: m(zz, noise_mode="const")[0], functional_only=True)
    img = torch.rand(2, 1, 16, 16, 16, generator=torch.Generator().manual_seed(7)) * 2 - 1
    same_as_stock(lambda: S.Discriminator(p, img_resolution=16), [img], dev, tol, functional_only=True)
    import b200conv3d as B
    n0 = B._lib.lib().b2c_launch_count()
    with installed(), torch.no_grad():
        y = S._upfirdn3d_ref(img.to(dev).repeat(1, 8, 1, 1, 1), torch.ones(4, 4, 4, device=dev) / 64, padding=2)
    assert tuple(y.shape) == (2, 8, 17, 17, 17)
    if dev != "cpu":
        assert B._lib.lib().b2c_launch_count() > n0


def check_convert_prepare(dev, tol):
    """convert() / prepare() on models built from STOCK layers (spectral-norm parametrised ones included): class swap
    in place, parameters untouched, same results, state_dict interchangeable both ways."""
    import b200conv3d as B
    from torch.nn.utils.parametrizations import spectral_norm
    torch.manual_seed(8)

    def make():
        return nn.Sequential(spectral_norm(nn.Conv3d(4, 8, 3, 1, 1)), nn.LeakyReLU(0.1, inplace=True),
                             nn.Conv3d(8, 8, 4, 2, 1, bias=False), nn.BatchNorm3d(8), nn.ReLU(True),
                             nn.ConvTranspose3d(8, 4, 4, 2, 1), nn.Tanh())
    ref, ours = make(), make()
    ours.load_state_dict(ref.state_dict())
    keys = list(ours.state_dict())
    ptrs = [p.data_ptr() for p in ours.parameters()]
    B.prepare(ours)
    assert list(ours.state_dict()) == keys and [p.data_ptr() for p in ours.parameters()] == ptrs
    assert isinstance(ours[0], B.nn.Conv3d) and isinstance(ours[3], B.nn.BatchNorm3d) and isinstance(ours[5], B.nn.ConvTranspose3d)
    assert isinstance(ours[1], B.nn.FusedAway) and ours[0].fused_act == "lrelu" and ours[3].fused_act == "relu"
    ours = ours.to(dev)
    x = torch.randn(2, 4, 6, 6, 6)
    xo, xr = x.clone().to(dev).requires_grad_(True), x.clone().requires_grad_(True)
    yo, yr = ours(xo), ref(xr)
    g = torch.randn_like(yr)
    yo.backward(g.to(dev)), yr.backward(g)
    assert _rel(yo, yr) < tol and _rel(xo.grad, xr.grad) < tol
    for (k, a), b in zip(ours.named_parameters(), ref.parameters()):
        assert _rel(a.grad, b.grad, 1e-3) < 10 * tol, k
    for k, v in ours.state_dict().items():             # spectral-norm u / v advanced identically
        assert _rel(v.float(), ref.state_dict()[k].float(), 1e-6) < 10 * tol, k
    ref.load_state_dict({k: v.cpu() for k, v in ours.state_dict().items()})
    ours.load_state_dict(ref.state_dict())


def check_reference_trainer_golden(dev, tol):
    """trainer.py unmodified (`Trainer.__init__`, `D_step`, `G_step` with `torch.cuda.amp.autocast()` + `GradScaler`,
    :80-81,208-283) on B200 layers through the launcher's setup() reproduces tests/golden/dcgan_ref.npz, which the
    stock reference produced (oracle/gen_golden.py)."""
    import b200conv3d as B
    from b200conv3d import run as launcher
    gold = np.load(os.path.join(ROOT, "tests", "golden", "dcgan_ref.npz"))
    width, z, batch = int(gold["width"]), int(gold["z"]), int(gold["batch"])
    stub_modules()
    seq_forward = nn.Sequential.forward
    launcher.setup(amp="keep" if dev == "cpu" else "bf16", fuse=True)     # install() + lazy fuse (+ bf16 autocast on CUDA)
    try:
        T = ref_import("trainer")
        p = argparse.Namespace(
            niters=1, batch_size=batch, z_size=z, filterG=width, filterD=width, iterD=1, lrG=5e-5, lrD=1e-4,
            data_path="", ngpu=1, steps_per_log=10 ** 9, steps_per_img_log=10 ** 9, log_dir=tempfile.mkdtemp(),
            device=dev, load_params=False, dcgan=True, hybrid=False, stylegan2=False, stylegan=False, msl=False,
            sngan=False, sagan=False, biggan=False, hinge=False)
        torch.manual_seed(20261104)
        tr = T.Trainer([torch.zeros(1)] * 8, p)
        _layers_are_ours(tr.netG), _layers_are_ours(tr.netD)
        for k, v in tr.netG.state_dict().items():
            assert torch.equal(v.cpu(), torch.from_numpy(gold["g0." + k])), k      # same init under the reference's seed
        torch.manual_seed(5)
        zin = torch.randn(batch, z, 1, 1, 1)
        real = torch.rand(batch, 1, 128, 128, 128) * 2 - 1
        with torch.no_grad():
            fake0 = tr.netG(zin.to(dev))
            dreal0 = tr.netD(real.to(dev))
        assert isinstance(tr.netG.main[2], B.nn.FusedAway)                         # the launcher fused ConvT-BN-ReLU
        assert _rel(fake0[:, 0, ::8, ::8, ::8], torch.from_numpy(gold["fake0_sample"])) < tol
        assert _rel(dreal0.reshape(-1), torch.from_numpy(gold["dreal0"])) < tol
        for q in list(tr.netD.parameters()) + list(tr.netG.parameters()):
            q.requires_grad = False
        torch.manual_seed(77)
        tr.D_step(0, real.to(dev))
        torch.manual_seed(78)
        tr.G_step(0)
        if dev == "cpu":      # CPU RNG draws the reference's noise: losses and the Adam step are the golden's
            np.testing.assert_allclose(tr.D_losses, gold["D_losses"], rtol=20 * tol, atol=2 * tol)
            np.testing.assert_allclose(tr.G_losses, gold["G_losses"], rtol=20 * tol, atol=2 * tol)
            for k, v in tr.netG.state_dict().items():
                if v.dim() == 5:
                    assert _rel(v, torch.from_numpy(gold["g1." + k])) < 2e-3, k
        else:                 # CUDA RNG: other noise, bf16 - the step must run and stay finite
            assert np.isfinite(tr.D_losses).all() and np.isfinite(tr.G_losses).all()
        assert all(torch.isfinite(q).all() for q in tr.netG.parameters())
    finally:
        B.uninstall()
        nn.Sequential.forward = seq_forward


def check_restated_models_under_install(dev, tol):
    """GPU box (no reference tree): install() on the pinned restatements oracle/ref_*.py, which - like the reference -
    only say `nn.Conv3d(...)` / `F.interpolate(...)`."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ref_biggan
    import ref_dcgan
    z = torch.randn(2, 8, 1, 1, 1, generator=torch.Generator().manual_seed(2))
    g, _ = same_as_stock(lambda: ref_dcgan.Generator(8, 2), [z], dev, tol)
    with torch.no_grad(), installed():
        img = g(z.to(dev)).float().cpu()
    same_as_stock(lambda: ref_dcgan.Discriminator(2), [img], dev, tol)
    up = lambda: functools.partial(F.interpolate, scale_factor=2)
    gen = torch.Generator().manual_seed(3)
    same_as_stock(lambda: ref_biggan.GBlockDeep(16, 8, upsample=up()), [torch.randn(2, 16, 4, 4, 4, generator=gen)], dev, tol)
    same_as_stock(lambda: ref_biggan.DBlockDeep(8, 16, downsample=nn.AvgPool3d(2)),
                  [torch.randn(2, 8, 8, 8, 8, generator=gen)], dev, tol)


CPU_CHECKS = ["check_reference_dcgan", "check_reference_biggan_blocks", "check_reference_fid_resnet",
              "check_reference_stylegan", "check_convert_prepare", "check_reference_trainer_golden"]


# ------------------------------------------------------------------------------------------------ CPU (subprocess, host oracle)
@needs_ref
@pytest.mark.parametrize("check", CPU_CHECKS)
def test_host(check, oracle_path):
    code = '''
import sys
sys.path.insert(0, %r); sys.path.insert(0, %r)
from b200conv3d import _lib
_lib.use_library(%r, host=True)
import test_install as T
getattr(T, %r)("cpu", 1e-4)
print("HOST-OK")
''' % (os.path.join(ROOT, "3d-gan_b200"), os.path.join(ROOT, "tests"), oracle_path, check)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=1200,
                         env=dict(os.environ, PYTHONDONTWRITEBYTECODE="1"))
    assert "HOST-OK" in out.stdout, out.stdout[-2000:] + out.stderr[-4000:]


def test_launcher_runs_an_unmodified_script(oracle_path, tmp_path):
    """python -m b200conv3d.run script.py args...: the script sees its own argv and builds B200 layers."""
    script = tmp_path / "user_script.py"
    script.write_text("import sys, torch.nn as nn\n"
                      "m = nn.Sequential(nn.Conv3d(2, 4, 3, 1, 1), nn.LeakyReLU(0.2))\n"
                      "print('ARGS', sys.argv[1:], type(m[0]).__module__, __name__)\n")
    env = dict(os.environ, PYTHONPATH=os.path.join(ROOT, "3d-gan_b200"), PYTHONDONTWRITEBYTECODE="1")
    out = subprocess.run([sys.executable, "-m", "b200conv3d.run", str(script), "--dcgan=True", "--batch_size", "4"],
                         capture_output=True, text=True, timeout=600, env=env)
    assert "ARGS ['--dcgan=True', '--batch_size', '4'] b200conv3d.nn __main__" in out.stdout, out.stdout + out.stderr[-3000:]


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_gpu_install_on_restated_models():
    check_restated_models_under_install("cuda", 1e-4)


@pytest.mark.gpu
def test_gpu_convert_prepare():
    check_convert_prepare("cuda", 1e-4)


@pytest.mark.gpu
@needs_ref
@pytest.mark.parametrize("check", CPU_CHECKS[:4] + CPU_CHECKS[5:])
def test_gpu_reference_files(check):
    globals()[check]("cuda", 1e-4)
