"""The BigGAN-deep 3D generator / discriminator of the reference built from the B200
drop-in layers: the caller side of the hot path for BASELINE config 3
(`--biggan=True --hinge=True --filterG=64 --filterD=64`, run.sh:1).

Reference: utils.py:8-20 (snconv3d / snlinear), utils.py:22-45 (Attention),
utils.py:47-86 (GBlockDeep), utils.py:88-132 (DBlockDeep), biggan.py:8-63 (Generator),
biggan.py:65-119 (Discriminator).  Module / parameter / buffer names equal the
reference's so state_dicts interchange with checkpoints of the reference
(trainer.py:153-163).

What runs where:
  * every conv (1x1x1, 3x3x3, the 1-channel stem and head) -> b2c_conv3d_* through
    :class:`b200conv3d.nn.Conv3d`; spectral norm stays torch's parametrisation on `.weight`
    (a fresh weight tensor per forward, re-packed by the layer; utils.py:13);
  * BatchNorm3d + ReLU pairs (utils.py:68-84, biggan.py:40-41) -> one fused b2c_bn_act pass;
  * the pre-activation ReLUs of DBlockDeep (utils.py:120-125) -> b2c_act;
  * channel drop + nearest x2 (utils.py:76-80) -> b2c_upsample2x; AvgPool3d(2) -> b2c_avgpool2
    (ATen has no channels-last 3-D kernels for either: it would convert to NCDHW and back);
  * the attention core (utils.py:33-43) is ONE fused library call, scaled_dot_product_attention with scale 1, on
    views of the NDHWC projections; residual adds, the shortcut cat, the attention max-pools
    and the two linears stay ATen ops on NDHWC tensors.
"""
import functools

import torch
import torch.nn as tnn
import torch.nn.functional as F
from torch.nn.utils.parametrizations import spectral_norm

from . import functional as BF
from . import nn as BN


def snconv3d(in_channels, out_channels, kernel_size=3, stride=1, padding=1, dilation=1, bias=True, sngan=False):
    """utils.py:8-14 (the inverted `sngan` flag is the reference's)."""
    conv = BN.Conv3d(in_channels=in_channels, out_channels=out_channels, kernel_size=kernel_size, stride=stride,
                     padding=padding, dilation=dilation, bias=bias)
    return conv if sngan else spectral_norm(conv)


def snlinear(in_features, out_features, sngan=False):
    """utils.py:16-20."""
    lin = tnn.Linear(in_features=in_features, out_features=out_features)
    return lin if sngan else spectral_norm(lin)


def _bn_relu(c):
    bn = BN.BatchNorm3d(c)
    bn.fused_act, bn.fused_slope = "relu", 0.0
    return bn


class Attention(tnn.Module):
    """utils.py:22-45.  The four 1x1x1 projections run on the B200 conv path."""

    def __init__(self, ch):
        super().__init__()
        self.ch, self.ch_ = ch, ch // 8
        self.f = snconv3d(ch, self.ch_, kernel_size=1, padding=0, bias=False)
        self.g = snconv3d(ch, self.ch_, kernel_size=1, padding=0, bias=False)
        self.h = snconv3d(ch, self.ch_, kernel_size=1, padding=0, bias=False)
        self.v = snconv3d(self.ch_, ch, kernel_size=1, padding=0, bias=False)
        self.gamma = tnn.Parameter(torch.tensor(0.), requires_grad=True)

    def forward(self, x, y=None):
        n, _, d, h, w = x.shape
        vox = d * h * w
        # utils.py:33-43 is single-head attention without the 1/sqrt(E) scale: queries = f(x) at every voxel, keys =
        # max-pooled g(x), values = max-pooled h(x).  softmax(q k^T) v goes through ONE fused library kernel
        # (scaled_dot_product_attention, scale = 1): the [n, vox, vox / 8] score tensor of the reference's two bmm +
        # softmax (8.6 GB in fp32 at the 32^3 site, batch 16) is never written.  In NDHWC memory the three
        # [n, voxels, channels] operands are views of the conv outputs.
        q = self.f(x).reshape(n, self.ch_, vox).permute(0, 2, 1)
        k = F.max_pool3d(self.g(x), [2, 2, 2], stride=2).reshape(n, self.ch_, vox // 8).permute(0, 2, 1)
        v = F.max_pool3d(self.h(x), [2, 2, 2], stride=2).reshape(n, self.ch_, vox // 8).permute(0, 2, 1)
        o = F.scaled_dot_product_attention(q.unsqueeze(1), k.unsqueeze(1), v.unsqueeze(1), scale=1.0).squeeze(1)
        o = self.v(o.permute(0, 2, 1).reshape(n, self.ch_, d, h, w))
        return self.gamma * o + x


class GBlockDeep(tnn.Module):
    """utils.py:47-86: BN-ReLU-1x1, BN-ReLU, [channel drop], [nearest x2], 3^3, BN-ReLU-3^3,
    BN-ReLU-1x1, + skip."""

    def __init__(self, in_channels, out_channels, upsample=None, sngan=False, channel_ratio=4):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.hidden_channels = hid = in_channels // channel_ratio
        self.conv1 = snconv3d(in_channels, hid, kernel_size=1, padding=0, sngan=sngan)
        self.conv2 = snconv3d(hid, hid, sngan=sngan)
        self.conv3 = snconv3d(hid, hid, sngan=sngan)
        self.conv4 = snconv3d(hid, out_channels, kernel_size=1, padding=0, sngan=sngan)
        self.bn1, self.bn2, self.bn3, self.bn4 = _bn_relu(in_channels), _bn_relu(hid), _bn_relu(hid), _bn_relu(hid)
        self.upsample = upsample
        self.activation = BN.FusedAway(tnn.ReLU(inplace=True))     # folded into the four BN passes
        # conv1 -> bn2 (utils.py:74-75, before the up-sampling), conv2 -> bn3, conv3 -> bn4 (utils.py:81-84) are
        # adjacent: their BatchNorm statistics ride in the conv's GEMM epilogue
        for conv in (self.conv1, self.conv2, self.conv3):
            conv.fused_stats = "channel"

    def forward(self, x):
        h = self.conv1(self.bn1(x))
        h = self.bn2(h)
        if self.upsample:          # utils.py:76-80: channel drop + nearest x2 of h and x, one pass each
            h = BF.upsample_nearest2x(h)
            x = BF.upsample_nearest2x(x, self.out_channels)
        elif self.in_channels != self.out_channels:
            x = x[:, :self.out_channels]
        h = self.conv2(h)
        h = self.conv3(self.bn3(h))
        h = self.conv4(self.bn4(h))
        return h + x


class DBlockDeep(tnn.Module):
    """utils.py:88-132: ReLU-1x1, ReLU-3^3, ReLU-3^3, ReLU, [AvgPool3d(2)], 1x1, + shortcut
    (pool, then cat[x, 1x1(x)] when channels grow)."""

    def __init__(self, in_channels, out_channels, wide=True, preactivation=True, downsample=None, channel_ratio=4):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.hidden_channels = hid = out_channels // channel_ratio
        self.preactivation = preactivation
        self.activation = BN.ReLU(inplace=True)
        self.downsample = downsample
        self.conv1 = snconv3d(in_channels, hid, kernel_size=1, padding=0)
        self.conv2 = snconv3d(hid, hid)
        self.conv3 = snconv3d(hid, hid)
        self.conv4 = snconv3d(hid, out_channels, kernel_size=1, padding=0)
        self.learnable_sc = in_channels != out_channels
        if self.learnable_sc:
            self.conv_sc = snconv3d(in_channels, out_channels - in_channels, kernel_size=1, padding=0)
        # the ReLUs after conv1 / conv2 / conv3 (utils.py:122-125) ride in the conv epilogues
        for c in (self.conv1, self.conv2, self.conv3):
            c.fused_act, c.fused_slope = "relu", 0.0

    def shortcut(self, x):
        if self.downsample:
            x = self.downsample(x)
        if self.learnable_sc:
            x = torch.cat([x, self.conv_sc(x)], 1)
        return x

    def forward(self, x):
        h = self.conv1(BF.activation(x, "relu"))     # F.relu(x), not in place: x feeds the shortcut
        h = self.conv2(h)
        h = self.conv3(h)
        if self.downsample:
            h = self.downsample(h)
        h = self.conv4(h)
        return h + self.shortcut(x)


def _orthogonal_init(model):
    """biggan.py:46-53 / 103-110.  On spectrally-normalised layers `module.weight` is the
    parametrised (derived) tensor, so the call only advances the RNG and the power iteration -
    reproduced as is so that seeds give the reference's initial state."""
    for m in model.modules():
        if isinstance(m, (tnn.Conv3d, tnn.Linear, tnn.Embedding)):
            tnn.init.orthogonal_(m.weight)


class Generator(tnn.Module):
    def __init__(self, z_size=128, filterG=64, biggan=True, sagan=False, sngan=False):
        super().__init__()
        self.dim_z = z_size
        f = filterG
        cin = [16 * f, 16 * f, 8 * f, 4 * f, 2 * f]
        cout = [16 * f, 8 * f, 4 * f, 2 * f, f]
        res = [8, 16, 32, 64, 128]
        up = functools.partial(F.interpolate, scale_factor=2)
        self.linear = snlinear(z_size, cin[0] * 4 ** 3, sngan=sngan)
        stages = []                       # biggan.py:23-27: one single-block list per g_index
        for i in range(5):
            if biggan:
                stages += [[GBlockDeep(cin[i], cin[i], upsample=None)], [GBlockDeep(cin[i], cout[i], upsample=up)]]
            else:
                stages += [[GBlockDeep(cin[i], cout[i], upsample=up, sngan=sngan)]]
            if (sagan or biggan) and res[i] == 32:
                stages[-1].append(Attention(cout[i]))
        self.blocks = tnn.ModuleList([tnn.ModuleList(s) for s in stages])
        head = snconv3d(cout[-1], 1, sngan=sngan)
        head.fused_act, head.fused_slope = "tanh", 0.0            # torch.tanh of biggan.py:63
        self.output_layer = tnn.Sequential(_bn_relu(cout[-1]), BN.FusedAway(tnn.ReLU(inplace=True)), head)
        _orthogonal_init(self)

    def forward(self, z):
        h = self.linear(z.squeeze())
        h = h.view(h.size(0), -1, 4, 4, 4)
        for stage in self.blocks:
            for block in stage:
                h = block(h)
        return self.output_layer(h)


class Discriminator(tnn.Module):
    def __init__(self, filterD=64, biggan=True, sagan=False, sngan=False):
        super().__init__()
        f = filterD
        cin = [f, 2 * f, 4 * f, 8 * f, 16 * f]
        cout = [2 * f, 4 * f, 8 * f, 16 * f, 16 * f]
        res = [64, 32, 16, 8, 4]
        self.input_conv = snconv3d(1, cin[0], sngan=sngan)
        stages = []
        for i in range(5):
            if biggan:
                stage = [DBlockDeep(cin[i], cout[i], downsample=BN.AvgPool3d(2)), DBlockDeep(cout[i], cout[i])]
            else:
                stage = [DBlockDeep(cin[i], cout[i], downsample=BN.AvgPool3d(2))]
            if (sagan or biggan) and res[i] == 16:
                stage.append(Attention(cout[i]))
            stages.append(tnn.ModuleList(stage))
        self.blocks = tnn.ModuleList(stages)
        self.linear = snlinear(cout[-1], 1, sngan=sngan)
        self.activation = BN.ReLU(inplace=True)
        _orthogonal_init(self)

    def forward(self, x):
        h = self.input_conv(x)
        for stage in self.blocks:
            for block in stage:
                h = block(h)
        h = torch.sum(self.activation(h), [2, 3, 4])
        return self.linear(h)


def build(z_size=128, filterG=64, filterD=64, device="cuda", **kw):
    g = Generator(z_size, filterG, **kw).to(device)
    d = Discriminator(filterD, **kw).to(device)
    return g, d
