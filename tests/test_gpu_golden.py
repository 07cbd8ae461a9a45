"""GPU parity of the whole DCGAN path against (1) the golden vectors of the UNMODIFIED
reference (tests/golden/dcgan_ref.npz, width 2: exact fp32 kernels, tol 1e-4) and (2) the
pinned CPU restatement oracle/ref_dcgan.py at a width that takes the tcgen05 path in bf16."""
import os
import sys

import numpy as np
import pytest
import torch

import b200conv3d as B
from abi import rel_err

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ref_dcgan  # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = os.path.join(ROOT, "tests", "golden", "dcgan_ref.npz")


def _load(net, gold, prefix):
    sd = net.state_dict()
    for k in list(sd):
        if prefix + k in gold.files:
            sd[k] = torch.from_numpy(gold[prefix + k])
    net.load_state_dict(sd)


def test_reference_golden_fp32():
    gold = np.load(GOLD)
    width, z, b = int(gold["width"]), int(gold["z"]), int(gold["batch"])
    g, d = B.dcgan.build(z, width, width, 128, "cpu", fuse=False)
    _load(g, gold, "g0.")
    _load(d, gold, "d0.")
    g, d = g.cuda(), d.cuda()
    B.fuse(g), B.fuse(d)
    torch.manual_seed(5)
    zin = torch.randn(b, z, 1, 1, 1)
    real = torch.rand(b, 1, 128, 128, 128) * 2 - 1
    with torch.no_grad():
        fake = g(zin.cuda())
        dreal = d(real.cuda())
    assert rel_err(fake[:, 0, ::8, ::8, ::8], torch.from_numpy(gold["fake0_sample"])) < 1e-4
    assert rel_err(dreal.reshape(-1), torch.from_numpy(gold["dreal0"])) < 1e-4
    # trainer.py:208-283, one D_step + G_step with the reference's noise draws
    step = B.train.GanStep(g, d, z, autocast_dtype=None)
    assert step.batch_real_fake        # LayerNorm D: D(real), D(fake) run as ONE batched pass - checked against the golden
    step.noise = lambda n: torch.randn(n, z, 1, 1, 1).cuda()
    torch.manual_seed(77)
    er, ef = step.D_step(real.cuda())
    torch.manual_seed(78)
    eg, _ = step.G_step(b)
    np.testing.assert_allclose([er.item(), ef.item()], gold["D_losses"][0], rtol=2e-4, atol=2e-5)
    np.testing.assert_allclose([eg.item()], gold["G_losses"], rtol=2e-4, atol=2e-5)
    for k, v in g.state_dict().items():
        if v.dim() == 5:
            assert rel_err(v, torch.from_numpy(gold["g1." + k])) < 2e-3, k     # one Adam step (sign-like update)


@pytest.mark.parametrize("width", [16])
def test_tensor_core_model_vs_oracle_bf16(width):
    """Width 16 (channels 16..256): every body layer takes the tcgen05 path, the 1-channel
    stem / head the thin path.  bf16 vs the fp32 CPU oracle on the same weights."""
    z, b = 32, 2
    torch.manual_seed(11)
    rg, rd = ref_dcgan.Generator(z, width), ref_dcgan.Discriminator(width)
    g, d = B.dcgan.build(z, width, width, 128, "cpu", fuse=False)
    g.load_state_dict(rg.state_dict()), d.load_state_dict(rd.state_dict())
    g, d = g.cuda(), d.cuda()
    B.fuse(g), B.fuse(d)
    zin = torch.randn(b, z, 1, 1, 1)
    real = torch.rand(b, 1, 128, 128, 128) * 2 - 1
    with torch.no_grad():
        f_ref, d_ref = rg(zin), rd(real)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            f, dd = g(zin.cuda()), d(real.cuda())
    assert B._lib.lib().b2c_last_algo() in (1, 2)
    ef = ((f.float().cpu() - f_ref).norm() / f_ref.norm()).item()
    ed = ((dd.float().cpu() - d_ref).abs().max() / d_ref.abs().max()).item()
    assert ef < 4e-2, ef          # 6 stacked bf16 layers with batch-2 BatchNorm: relative L2
    assert ed < 4e-2, ed


def test_fid_resnet_vs_reference_golden():
    """FID_ResNet.resnet50 features (eval_utils.py:41-44) on the product layers, fp32 (exact
    kernels, 1e-3 through 53 train-mode BatchNorms) and bf16 (tcgen05 path)."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "fid_resnet_ref.npz"))
    import ref_fid_resnet
    torch.manual_seed(int(g["w_seed"]))
    ref = ref_fid_resnet.resnet50()
    m = B.fid_resnet.resnet50()
    m.load_state_dict(ref.state_dict())
    m = m.cuda()
    torch.manual_seed(int(g["x_seed"]))
    x = torch.randn(2, 1, 32, 32, 32).cuda()
    with torch.no_grad():
        y = m(x)
    assert tuple(y.shape) == tuple(g["y_shape"])
    assert rel_err(y[:, ::64], torch.from_numpy(g["y_sample"])) < 1e-3
    f32 = B.fid_resnet.features(m, x)
    assert rel_err(f32, torch.from_numpy(g["feat"])) < 1e-3
    with torch.autocast("cuda", dtype=torch.bfloat16):
        fbf = B.fid_resnet.features(m, x)
    ref_feat = torch.from_numpy(g["feat"])
    l2 = ((fbf.cpu() - ref_feat).norm() / ref_feat.norm()).item()
    # 53 stacked bf16 layers with 2-sample train-mode BatchNorm: bounded by what stock torch
    # bf16 autocast (cuDNN) loses on the same GPU with the same weights
    wit = ref.cuda()
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        fw = wit(x).float().mean(dim=(2, 3, 4))
    l2w = ((fw.cpu() - ref_feat).norm() / ref_feat.norm()).item()
    assert l2 < max(4e-2, 1.5 * l2w), (l2, l2w)


# ---------------------------------------------------------------- BigGAN-deep (BASELINE config 3)
def _biggan_from_gold(gold):
    import ref_biggan  # noqa: F401
    width, z = int(gold["width"]), int(gold["z"])
    g, d = B.biggan.build(z, width, width, "cpu")
    g.load_state_dict({k: torch.from_numpy(gold["g0." + k]) for k in g.state_dict()})
    d.load_state_dict({k: torch.from_numpy(gold["d0." + k]) for k in d.state_dict()})
    return g.cuda(), d.cuda(), z


def test_biggan_reference_golden_fp32():
    """Product BigGAN-deep (b200conv3d.biggan) against golden vectors of the UNMODIFIED reference
    (biggan.py / utils.py / trainer.py hinge branch), fp32 exact kernels: eval-mode outputs, then
    one Trainer.D_step + G_step with the reference's noise draws - losses, every parameter, the
    spectral-norm u / v buffers and BN running statistics."""
    gold = np.load(os.path.join(ROOT, "tests", "golden", "biggan_ref.npz"))
    g, d, z = _biggan_from_gold(gold)
    b = int(gold["batch"])
    assert sorted("g0." + k for k in g.state_dict()) == sorted(k for k in gold.files if k.startswith("g0."))
    assert sorted("d0." + k for k in d.state_dict()) == sorted(k for k in gold.files if k.startswith("d0."))
    torch.manual_seed(5)
    zin = torch.randn(b, z, 1, 1, 1)
    real = torch.rand(b, 1, 128, 128, 128) * 2 - 1
    g.eval(), d.eval()
    with torch.no_grad():
        fake = g(zin.cuda())
        dreal = d(real.cuda())
    g.train(), d.train()
    assert rel_err(fake[:, 0, ::8, ::8, ::8], torch.from_numpy(gold["fake0_sample"])) < 1e-4
    assert rel_err(dreal.reshape(-1), torch.from_numpy(gold["dreal0"])) < 1e-4
    step = B.train.GanStep(g, d, z, hinge=True, autocast_dtype=None)
    assert not step.batch_real_fake    # spectral norm: one power iteration per forward, the two passes must stay two
    step.noise = lambda n: torch.randn(n, z, 1, 1, 1).cuda()
    torch.manual_seed(77)
    er, ef = step.D_step(real.cuda())
    torch.manual_seed(78)
    eg, _ = step.G_step(b)
    np.testing.assert_allclose([er.item(), ef.item()], gold["D_losses"][0], rtol=2e-4, atol=2e-5)
    np.testing.assert_allclose([eg.item()], gold["G_losses"], rtol=2e-4, atol=2e-5)
    for pre, m in (("g1.", g), ("d1.", d)):
        for k, v in m.state_dict().items():
            ref = torch.from_numpy(gold[pre + k])
            if "running" in k or k.endswith("._u") or k.endswith("._v"):
                assert rel_err(v, ref) < 1e-3, k
            elif v.dim() >= 1 and v.numel() > 1:
                # one Adam step is sign-like (lr * g / |g|): entries whose gradient is ~0 may flip
                bad = ((v.cpu() - ref).abs() > 2e-3 * ref.abs().max() + 1e-6).float().mean().item()
                assert bad < max(0.02, 1.5 / v.numel()), (k, bad)


def test_biggan_tensor_core_vs_oracle_bf16():
    """Width 32 (hidden channels 8..128, stem / head 1 channel): bf16 through the tcgen05 / thin
    paths against the fp32 CPU oracle on the same weights, forward and parameter gradients of one
    hinge discriminator loss."""
    import ref_biggan
    width, z, b = 32, 16, 2
    torch.manual_seed(21)
    rd, rg = ref_biggan.Discriminator(width), ref_biggan.Generator(z, width)
    g, d = B.biggan.build(z, width, width, "cpu")
    g.load_state_dict(rg.state_dict()), d.load_state_dict(rd.state_dict())
    g, d = g.cuda(), d.cuda()
    zin = torch.randn(b, z, 1, 1, 1)
    real = torch.rand(b, 1, 128, 128, 128) * 2 - 1
    rg.eval(), rd.eval(), g.eval(), d.eval()      # frozen spectral norm / BN statistics: same function
    with torch.no_grad():
        f_ref, d_ref = rg(zin), rd(real)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            f, dd = g(zin.cuda()), d(real.cuda())
        # stock torch layers under bf16 autocast on the same GPU: the bf16 noise floor (witness)
        wg, wd = rg.cuda(), rd.cuda()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            fw, dw = wg(zin.cuda()), wd(real.cuda())
    ef = ((f.float().cpu() - f_ref).norm() / f_ref.norm()).item()
    efw = ((fw.float().cpu() - f_ref).norm() / f_ref.norm()).item()
    ed = ((dd.float().cpu() - d_ref).abs().max() / d_ref.abs().max()).item()
    edw = ((dw.float().cpu() - d_ref).abs().max() / d_ref.abs().max()).item()
    assert ef < max(4e-2, 1.5 * efw), (ef, efw)
    assert ed < max(4e-2, 1.5 * edw), (ed, edw)
