"""The DCGAN generator / discriminator layer tables of the reference, built from the
B200 drop-in layers.  This is the caller side of the hot path for BASELINE config 2
(reference: dcgan.py:45-70 Generator.main, dcgan.py:158-182 Discriminator.main default
branch, init dcgan.py:74-81).  When /root/reference is importable the same models are
obtained by ``b200conv3d.install(); import dcgan`` — this module exists so that bench.py
and the GPU tests do not need the reference tree on the GPU box.  Module / parameter
names equal the reference's (``main.<idx>.weight``), so state_dicts interchange.
"""
import torch
import torch.nn as tnn

from . import nn as BN
from . import patch


def _g_layers(m, nz, ngf, nc):
    L = []
    # dcgan.py:47-49  z x 1^3 -> (16 ngf) x 4^3
    L += [m.ConvTranspose3d(nz, ngf * 16, 4, stride=1), m.BatchNorm3d(ngf * 16), m.ReLU(True)]
    # dcgan.py:51-65  four stride-2 stages, 4^3 -> 64^3
    for cin, cout in ((16, 8), (8, 4), (4, 2), (2, 1)):
        L += [m.ConvTranspose3d(ngf * cin, ngf * cout, 4, stride=2, padding=1),
              m.BatchNorm3d(ngf * cout), m.ReLU(True)]
    # dcgan.py:67-68  ngf x 64^3 -> 1 x 128^3, Tanh
    L += [m.ConvTranspose3d(ngf, nc, 4, stride=2, padding=1), m.Tanh()]
    return L


def _d_layers(m, ndf, nc, res):
    L = []
    # dcgan.py:161-179  five stride-2 stages with LayerNorm([C, r, r, r]) + LeakyReLU(0.2)
    cin, r = nc, res
    for mult in (1, 2, 4, 8, 16):
        r //= 2
        L += [m.Conv3d(cin, ndf * mult, 4, stride=2, padding=1, bias=False),
              m.LayerNorm([ndf * mult, r, r, r]), m.LeakyReLU(0.2, inplace=True)]
        cin = ndf * mult
    # dcgan.py:181  (16 ndf) x 4^3 -> 1
    L += [m.Conv3d(cin, 1, (r, r, r), stride=1, padding=0, bias=False)]
    return L


def weights_init(mod):
    """dcgan.py:74-81 / 184-191: conv ~ N(0, 0.02); BatchNorm gamma ~ N(1, 0.02), beta 0."""
    name = mod.__class__.__name__
    if name.find("Conv") != -1:
        tnn.init.normal_(mod.weight.data, 0.0, 0.02)
    elif name.find("BatchNorm") != -1:
        tnn.init.normal_(mod.weight.data, 1.0, 0.02)
        tnn.init.constant_(mod.bias.data, 0)


class Generator(tnn.Module):
    def __init__(self, z_size=512, filterG=128, layers=BN):
        super().__init__()
        self.dim_z = z_size
        self.main = tnn.Sequential(*_g_layers(layers, z_size, filterG, 1))
        self.apply(weights_init)

    def forward(self, z):
        return self.main(z)


class Discriminator(tnn.Module):
    def __init__(self, filterD=128, res=128, layers=BN):
        super().__init__()
        self.main = tnn.Sequential(*_d_layers(layers, filterD, 1, res))
        self.apply(weights_init)

    def forward(self, x):
        return self.main(x)


def build(z_size=512, filterG=128, filterD=128, res=128, device="cuda", fuse=True):
    """Both nets on `device`, activations folded into their producers."""
    assert res == 128, "the reference generator always emits 128^3 (dcgan.py:47-69)"
    g = Generator(z_size, filterG).to(device)
    d = Discriminator(filterD, res).to(device)
    if fuse:
        patch.fuse(g)
        patch.fuse(d)
    return g, d
