"""Pins oracle/ref_dcgan.py (the CPU restatement used as bench.py's reference arm and
as the module-level oracle) against golden vectors produced by the UNMODIFIED reference
(/root/reference dcgan.py + trainer.py, see oracle/gen_golden.py)."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ref_dcgan  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden", "dcgan_ref.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def _build(gold):
    width, z = int(gold["width"]), int(gold["z"])
    g, d = ref_dcgan.Generator(z, width), ref_dcgan.Discriminator(width)
    sd = g.state_dict()
    for k in sd:
        sd[k] = torch.from_numpy(gold["g0." + k])
    g.load_state_dict(sd)
    sd = d.state_dict()
    for k in list(sd):
        if "d0." + k in gold.files:
            sd[k] = torch.from_numpy(gold["d0." + k])
    d.load_state_dict(sd)
    return g, d, z


def _inputs(gold):
    torch.manual_seed(5)
    b, z = int(gold["batch"]), int(gold["z"])
    zin = torch.randn(b, z, 1, 1, 1)
    real = torch.rand(b, 1, 128, 128, 128) * 2 - 1
    np.testing.assert_array_equal(zin.numpy(), gold["zin"])
    np.testing.assert_array_equal(real[:, 0, ::16, ::16, ::16].numpy(), gold["real_sample"])
    return zin, real


def test_state_dict_keys_match_reference(gold):
    g, d, _ = _build(gold)
    assert sorted("g0." + k for k in g.state_dict()) == sorted(k for k in gold.files if k.startswith("g0."))
    ref_conv = sorted(k for k in gold.files if k.startswith("d0."))
    ours_conv = sorted("d0." + k for k, v in d.state_dict().items() if v.dim() == 5 and v.shape[2] == 4)
    assert ref_conv == ours_conv


def test_forward_matches_reference(gold):
    g, d, _ = _build(gold)
    zin, real = _inputs(gold)
    with torch.no_grad():
        fake = g(zin)
        dreal = d(real)
    np.testing.assert_allclose(fake[:, 0, ::8, ::8, ::8].numpy(), gold["fake0_sample"], rtol=1e-5, atol=1e-7)
    assert abs(fake.double().sum().item() - float(gold["fake0_sum"])) < 1e-6 * fake.numel() ** 0.5 + 1e-3
    np.testing.assert_allclose(dreal.reshape(-1).numpy(), gold["dreal0"], rtol=1e-5, atol=1e-7)


def test_train_phases_match_reference(gold):
    """One Trainer.D_step + G_step (trainer.py:208-283): losses and updated filters."""
    g, d, z = _build(gold)
    _, real = _inputs(gold)
    tr = ref_dcgan.RefTrainer(g, d, z)
    torch.manual_seed(77)
    er, ef = tr.d_step(real)
    torch.manual_seed(78)
    eg, _ = tr.g_step(int(gold["batch"]))
    np.testing.assert_allclose([er, ef], gold["D_losses"][0], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose([eg], gold["G_losses"], rtol=1e-5, atol=1e-7)
    for k, v in g.state_dict().items():
        if v.dim() == 5:
            np.testing.assert_allclose(v.numpy(), gold["g1." + k], rtol=1e-4, atol=1e-6)
    for k, v in d.state_dict().items():
        if v.dim() == 5 and v.shape[2] == 4:
            np.testing.assert_allclose(v.numpy(), gold["d1." + k], rtol=1e-4, atol=1e-6)
    sd = d.state_dict()
    np.testing.assert_allclose(sd["main.1.weight"][:, ::16, ::16, ::16].numpy(), gold["d1.ln1_weight_sample"],
                               rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(sd["main.1.bias"][:, ::16, ::16, ::16].numpy(), gold["d1.ln1_bias_sample"],
                               rtol=1e-4, atol=1e-6)


def test_fid_resnet_port_matches_reference():
    """oracle/ref_fid_resnet.py vs golden of the unmodified FID_ResNet.resnet50 (seed replay)."""
    import ref_fid_resnet
    g = np.load(os.path.join(ROOT, "tests", "golden", "fid_resnet_ref.npz"))
    torch.manual_seed(int(g["w_seed"]))
    m = ref_fid_resnet.resnet50()
    sd = m.state_dict()
    assert abs(sd["conv1.weight"].double().sum().item() - float(g["conv1_w_sum"])) < 1e-9
    assert abs(sd["layer4.2.conv3.weight"].double().sum().item() - float(g["l4_w_sum"])) < 1e-9
    torch.manual_seed(int(g["x_seed"]))
    x = torch.randn(2, 1, 32, 32, 32)
    with torch.no_grad():
        y = m(x)
    assert tuple(y.shape) == tuple(g["y_shape"])
    np.testing.assert_allclose(y[:, ::64].numpy(), g["y_sample"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(y.mean(dim=(2, 3, 4)).numpy(), g["feat"], rtol=1e-4, atol=1e-5)


# ---------------------------------------------------------------- BigGAN-deep (BASELINE config 3)
import ref_biggan  # noqa: E402

GOLD_BIG = os.path.join(ROOT, "tests", "golden", "biggan_ref.npz")


@pytest.fixture(scope="module")
def gold_big():
    return np.load(GOLD_BIG)


def _build_big(gold, load=True):
    width, z = int(gold["width"]), int(gold["z"])
    torch.manual_seed(20261104)
    d = ref_biggan.Discriminator(width)            # trainer.py:66-67 builds D first
    g = ref_biggan.Generator(z, width)
    if load:
        g.load_state_dict({k: torch.from_numpy(gold["g0." + k]) for k in g.state_dict()})
        d.load_state_dict({k: torch.from_numpy(gold["d0." + k]) for k in d.state_dict()})
    return g, d, z


def test_biggan_init_and_names_match_reference(gold_big):
    """Same seed -> same state_dict, key for key: pins the spectral-norm wrapping, the
    (ineffective on SN layers) orthogonal init and the RNG consumption order (biggan.py:46-53)."""
    g, d, _ = _build_big(gold_big, load=False)
    for pre, m in (("g0.", g), ("d0.", d)):
        sd = m.state_dict()
        assert sorted(pre + k for k in sd) == sorted(k for k in gold_big.files if k.startswith(pre))
        for k, v in sd.items():
            np.testing.assert_array_equal(v.numpy(), gold_big[pre + k], err_msg=k)


def test_biggan_forward_matches_reference(gold_big):
    g, d, _ = _build_big(gold_big)
    zin, real = _inputs(gold_big)
    g.eval(), d.eval()
    with torch.no_grad():
        fake = g(zin)
        dreal = d(real)
    np.testing.assert_allclose(fake[:, 0, ::8, ::8, ::8].numpy(), gold_big["fake0_sample"], rtol=1e-5, atol=1e-7)
    assert abs(fake.double().sum().item() - float(gold_big["fake0_sum"])) < 1e-6 * fake.numel() ** 0.5 + 1e-3
    np.testing.assert_allclose(dreal.reshape(-1).numpy(), gold_big["dreal0"], rtol=1e-5, atol=1e-6)


def test_biggan_hinge_phases_match_reference(gold_big):
    """Trainer.D_step (hinge, trainer.py:222-233) + G_step: losses, every parameter and buffer
    (spectral-norm u / v, BN running statistics) after the two optimizer steps."""
    g, d, z = _build_big(gold_big)
    _, real = _inputs(gold_big)
    tr = ref_dcgan.RefTrainer(g, d, z, hinge=True)
    torch.manual_seed(77)
    er, ef = tr.d_step(real)
    torch.manual_seed(78)
    eg, _ = tr.g_step(int(gold_big["batch"]))
    np.testing.assert_allclose([er, ef], gold_big["D_losses"][0], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose([eg], gold_big["G_losses"], rtol=1e-5, atol=1e-7)
    for pre, m in (("g1.", g), ("d1.", d)):
        for k, v in m.state_dict().items():
            np.testing.assert_allclose(v.numpy(), gold_big[pre + k], rtol=2e-4, atol=2e-6, err_msg=k)


def test_batched_discriminator_pass_is_allow_listed():
    """GanStep runs D(real), D(fake) as ONE pass only for discriminators made of per-sample, stateless layers
    (trainer.py:228-241 semantics): anything with batch coupling, per-call state or RNG keeps the two passes."""
    import torch.nn as tnn
    import b200conv3d as B
    from b200conv3d.train import per_sample_stateless
    _, d = B.dcgan.build(16, 8, 8, 128, torch.device("cpu"))
    assert per_sample_stateless(d)
    _, d_sn = B.biggan.build(16, 8, 8, torch.device("cpu"))
    assert not per_sample_stateless(d_sn)                                             # spectral norm
    assert not per_sample_stateless(tnn.Sequential(tnn.Conv3d(1, 4, 3), tnn.BatchNorm3d(4)))
    assert not per_sample_stateless(tnn.Sequential(tnn.Conv3d(1, 4, 3), tnn.Dropout(0.1)))
    assert not per_sample_stateless(tnn.Sequential(tnn.Conv3d(1, 4, 3), tnn.RReLU()))  # unknown layer: not batched


def test_loss_curve_criterion_on_recorded_curves():
    """tools/loss_curves.py judges the recorded 100-step curves (profiles/) over the horizon where stock bf16 still
    tracks fp32; the recorded run must pass and a curve that leaves the envelope must fail."""
    import copy
    import json
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import loss_curves
    d = json.load(open(os.path.join(ROOT, "profiles", "r02_loss_curves_100steps.json")))
    res = loss_curves.evaluate(d["ref"], d["ours"], d["witness"])
    assert res["pass"] and 20 <= res["horizon_steps"] <= 100
    assert all(c["window_rel"] < 0.05 for c in res["curves"].values())
    bad = copy.deepcopy(d["ours"])
    k = next(iter(bad))
    bad[k] = [v + 0.2 * res["curves"][k]["window_range"] for v in bad[k]]        # a constant 20 % offset from step 0
    assert not loss_curves.evaluate(d["ref"], bad, d["witness"])["pass"]
