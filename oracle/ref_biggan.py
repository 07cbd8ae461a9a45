"""ORACLE (test infrastructure, never on the product path): CPU restatement of the reference's
BigGAN-deep 3D generator / discriminator from stock PyTorch layers (BASELINE config 3).

  sn_conv / sn_linear <- /root/reference/utils.py:8-20   (spectral_norm unless sngan=True: the
                          reference's flag is inverted, SURVEY.md §8(c) quirks - kept)
  Attention           <- /root/reference/utils.py:22-45
  GBlockDeep          <- /root/reference/utils.py:47-86
  DBlockDeep          <- /root/reference/utils.py:88-132
  Generator           <- /root/reference/biggan.py:8-63   (orthogonal init :46-53)
  Discriminator       <- /root/reference/biggan.py:65-119
  hinge D_step        <- /root/reference/trainer.py:222-233 (in ref_dcgan-style step helpers of the tests)

Module / parameter names equal the reference's, so state_dicts are interchangeable
(`blocks.<stage>.<j>.conv1.parametrizations.weight.original`, ...).

Pinned: tests/test_oracle_models.py checks this file against tests/golden/biggan_ref.npz, which
oracle/gen_golden_biggan.py produced from the UNMODIFIED reference modules in the build container.
Only tests/, smoke() and bench tools' CPU legs import this file.
"""
import functools

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.nn.utils.parametrizations import spectral_norm


def sn_conv(cin, cout, k=3, pad=1, bias=True, sngan=False):
    conv = nn.Conv3d(in_channels=cin, out_channels=cout, kernel_size=k, stride=1, padding=pad, dilation=1, bias=bias)
    return conv if sngan else spectral_norm(conv)


def sn_linear(fin, fout, sngan=False):
    lin = nn.Linear(in_features=fin, out_features=fout)
    return lin if sngan else spectral_norm(lin)


class Attention(nn.Module):
    def __init__(self, ch):
        super().__init__()
        self.ch, self.ch_ = ch, ch // 8
        self.f = sn_conv(ch, self.ch_, k=1, pad=0, bias=False)
        self.g = sn_conv(ch, self.ch_, k=1, pad=0, bias=False)
        self.h = sn_conv(ch, self.ch_, k=1, pad=0, bias=False)
        self.v = sn_conv(self.ch_, ch, k=1, pad=0, bias=False)
        self.gamma = nn.Parameter(torch.tensor(0.), requires_grad=True)

    def forward(self, x, y=None):
        vox = x.shape[2] * x.shape[3] * x.shape[4]
        f = self.f(x).view(-1, self.ch_, vox)
        g = F.max_pool3d(self.g(x), [2, 2, 2], stride=2).view(-1, self.ch_, vox // 8)
        beta = F.softmax(torch.bmm(f.permute(0, 2, 1), g), -1)
        h = F.max_pool3d(self.h(x), [2, 2, 2], stride=2).view(-1, self.ch_, vox // 8)
        o = self.v(torch.bmm(h, beta.permute(0, 2, 1)).view(-1, self.ch_, x.shape[2], x.shape[3], x.shape[4]))
        return self.gamma * o + x


class GBlockDeep(nn.Module):
    def __init__(self, in_channels, out_channels, upsample=None, sngan=False, channel_ratio=4):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        hid = self.hidden_channels = in_channels // channel_ratio
        self.conv1 = sn_conv(in_channels, hid, k=1, pad=0, sngan=sngan)
        self.conv2 = sn_conv(hid, hid, sngan=sngan)
        self.conv3 = sn_conv(hid, hid, sngan=sngan)
        self.conv4 = sn_conv(hid, out_channels, k=1, pad=0, sngan=sngan)
        self.bn1 = nn.BatchNorm3d(in_channels)
        self.bn2 = nn.BatchNorm3d(hid)
        self.bn3 = nn.BatchNorm3d(hid)
        self.bn4 = nn.BatchNorm3d(hid)
        self.upsample = upsample
        self.activation = nn.ReLU(inplace=True)

    def forward(self, x):
        h = self.conv1(self.activation(self.bn1(x)))
        h = self.activation(self.bn2(h))
        if self.in_channels != self.out_channels:
            x = x[:, :self.out_channels]
        if self.upsample:
            h, x = self.upsample(h), self.upsample(x)
        h = self.conv2(h)
        h = self.conv3(self.activation(self.bn3(h)))
        h = self.conv4(self.activation(self.bn4(h)))
        return h + x


class DBlockDeep(nn.Module):
    def __init__(self, in_channels, out_channels, wide=True, preactivation=True, downsample=None, channel_ratio=4):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        hid = self.hidden_channels = out_channels // channel_ratio
        self.preactivation = preactivation
        self.activation = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.conv1 = sn_conv(in_channels, hid, k=1, pad=0)
        self.conv2 = sn_conv(hid, hid)
        self.conv3 = sn_conv(hid, hid)
        self.conv4 = sn_conv(hid, out_channels, k=1, pad=0)
        self.learnable_sc = in_channels != out_channels
        if self.learnable_sc:
            self.conv_sc = sn_conv(in_channels, out_channels - in_channels, k=1, pad=0)

    def shortcut(self, x):
        if self.downsample:
            x = self.downsample(x)
        if self.learnable_sc:
            x = torch.cat([x, self.conv_sc(x)], 1)
        return x

    def forward(self, x):
        h = self.conv1(F.relu(x))
        h = self.conv2(self.activation(h))
        h = self.conv3(self.activation(h))
        h = self.activation(h)
        if self.downsample:
            h = self.downsample(h)
        h = self.conv4(h)
        return h + self.shortcut(x)


def _orthogonal(model):
    for m in model.modules():
        if isinstance(m, (nn.Conv3d, nn.Linear, nn.Embedding)):
            nn.init.orthogonal_(m.weight)


class Generator(nn.Module):
    def __init__(self, z_size=128, filterG=64, biggan=True, sagan=False, sngan=False):
        super().__init__()
        f = filterG
        cin = [16 * f, 16 * f, 8 * f, 4 * f, 2 * f]
        cout = [16 * f, 8 * f, 4 * f, 2 * f, f]
        res = [8, 16, 32, 64, 128]
        up = functools.partial(F.interpolate, scale_factor=2)
        self.linear = sn_linear(z_size, cin[0] * 4 ** 3, sngan=sngan)
        # biggan.py:23-27 appends one single-block list per g_index, so the generator's ModuleList
        # has TEN one-block entries (blocks.0.0, blocks.1.0, ...); attention joins the last of them
        stages = []
        for i in range(5):
            if biggan:
                stages += [[GBlockDeep(cin[i], cin[i], upsample=None)], [GBlockDeep(cin[i], cout[i], upsample=up)]]
            else:
                stages += [[GBlockDeep(cin[i], cout[i], upsample=up, sngan=sngan)]]
            if (sagan or biggan) and res[i] == 32:
                stages[-1].append(Attention(cout[i]))
        self.blocks = nn.ModuleList([nn.ModuleList(st) for st in stages])
        self.output_layer = nn.Sequential(nn.BatchNorm3d(cout[-1]), nn.ReLU(inplace=True), sn_conv(cout[-1], 1, sngan=sngan))
        _orthogonal(self)

    def forward(self, z):
        h = self.linear(z.squeeze())
        h = h.view(h.size(0), -1, 4, 4, 4)
        for stage in self.blocks:
            for block in stage:
                h = block(h)
        return torch.tanh(self.output_layer(h))


class Discriminator(nn.Module):
    def __init__(self, filterD=64, biggan=True, sagan=False, sngan=False):
        super().__init__()
        f = filterD
        cin = [f, 2 * f, 4 * f, 8 * f, 16 * f]
        cout = [2 * f, 4 * f, 8 * f, 16 * f, 16 * f]
        res = [64, 32, 16, 8, 4]
        self.input_conv = sn_conv(1, cin[0], sngan=sngan)
        stages = []
        for i in range(5):
            if biggan:
                stage = [DBlockDeep(cin[i], cout[i], downsample=nn.AvgPool3d(2)), DBlockDeep(cout[i], cout[i], downsample=None)]
            else:
                stage = [DBlockDeep(cin[i], cout[i], downsample=nn.AvgPool3d(2))]
            if (sagan or biggan) and res[i] == 16:
                stage.append(Attention(cout[i]))
            stages.append(nn.ModuleList(stage))
        self.blocks = nn.ModuleList(stages)
        self.linear = sn_linear(cout[-1], 1, sngan=sngan)
        self.activation = nn.ReLU(inplace=True)
        _orthogonal(self)

    def forward(self, x):
        h = self.input_conv(x)
        for stage in self.blocks:
            for block in stage:
                h = block(h)
        h = torch.sum(self.activation(h), [2, 3, 4])
        return self.linear(h)
