"""The StyleGAN networks of the reference on the B200 path: the caller side of the hot path for
BASELINE config 4.

  FullyConnectedLayer / MappingNetwork       <- stylegan.py:298-392
  Conv3dLayer                                <- stylegan.py:446-487
  SynthesisLayer / OutBlock / SynthesisBlock <- stylegan.py:489-640
  SynthesisNetwork / Generator               <- stylegan.py:644-718
  DiscriminatorBlock / MinibatchStdLayer / DiscriminatorEpilogue / Discriminator
                                             <- stylegan.py:721-928
  AdaIN / styleGAN_gen (the v1-like generator) <- stylegan.py:931-1149
  StyleGAN2Loss (step_D / step_G, R1 and path-length terms) <- stylegan.py:6-100

Module / parameter / buffer names and the order in which the constructors draw from the RNG
equal the reference's: state_dicts interchange and a model built under the reference's seed is
bit-identical (tests/test_stylegan_net.py against tests/golden/stylegan_net_ref.npz).

What runs where: every convolution goes through :mod:`b200conv3d.stylegan_ops`
(`modulated_conv3d` / `conv3d_resample` -> b2c_conv3d_* tcgen05 kernels, FIR resampling ->
b2c_fir3d); bias + activation -> b2c_act; the small dense layers (mapping network, style
affines, the discriminator's two FC layers) are ATen GEMMs (library GEMMs, not on the conv
path); minibatch-std, instance-norm statistics of AdaIN and trilinear x2 stay ATen reductions.

Kept quirks of the reference: `gain` arguments of the layers are accepted
and ignored (stylegan.py:478,512), the skip-image FIR uses gain up**3 while the conv FIR uses
up**2, `noise_const` is 2-D [R, R] and broadcasts over the two last axes, R1 sums over
dims [1,2,3] only, `MinibatchStdLayer` names the axes H, W, D.
"""
import math

import numpy as np
import torch
import torch.nn as tnn
import torch.nn.functional as F

from . import functional as BF
from . import nn as BN
from . import stylegan_ops as SO


def normalize_2nd_moment(x, dim=1, eps=1e-8):
    """stylegan.py:115-116."""
    return x * (x.square().mean(dim=dim, keepdim=True) + eps).rsqrt()


class FullyConnectedLayer(tnn.Module):
    """stylegan.py:298-327: equalised-learning-rate dense layer (run-time weight gain)."""

    def __init__(self, in_features, out_features, bias=True, activation="linear", lr_multiplier=1, bias_init=0):
        super().__init__()
        self.activation = activation
        self.weight = tnn.Parameter(torch.randn([out_features, in_features]) / lr_multiplier)
        self.bias = tnn.Parameter(torch.full([out_features], np.float32(bias_init))) if bias else None
        self.weight_gain = lr_multiplier / np.sqrt(in_features)
        self.bias_gain = lr_multiplier

    def forward(self, x):
        w = self.weight.to(x.dtype) * self.weight_gain
        b = self.bias
        if b is not None:
            b = b.to(x.dtype)
            if self.bias_gain != 1:
                b = b * self.bias_gain
        if self.activation == "linear" and b is not None:
            return torch.addmm(b.unsqueeze(0), x, w.t())
        x = x.matmul(w.t())
        if b is not None:
            x = x + b
        if self.activation == "lrelu":
            return F.leaky_relu(x, 0.2)
        if self.activation == "linear":
            return x
        raise NotImplementedError("activation %r is not used by the reference's dense layers" % (self.activation,))


class MappingNetwork(tnn.Module):
    """stylegan.py:329-392 (no conditioning label; z is squeezed, normalised, mapped, broadcast)."""

    def __init__(self, z_dim, w_dim, num_ws, num_layers=8, layer_features=None, activation="lrelu",
                 lr_multiplier=0.01, w_avg_beta=0.995):
        super().__init__()
        self.z_dim, self.w_dim, self.num_ws = z_dim, w_dim, num_ws
        self.num_layers, self.w_avg_beta = num_layers, w_avg_beta
        feats = [z_dim] + [w_dim if layer_features is None else layer_features] * (num_layers - 1) + [w_dim]
        for i in range(num_layers):
            setattr(self, "fc%d" % i, FullyConnectedLayer(feats[i], feats[i + 1], activation=activation,
                                                          lr_multiplier=lr_multiplier))
        if num_ws is not None and w_avg_beta is not None:
            self.register_buffer("w_avg", torch.zeros([w_dim]))

    def forward(self, z, truncation_psi=1, truncation_cutoff=None, skip_w_avg_update=True):
        x = None
        z = z.squeeze()
        if self.z_dim > 0:
            x = normalize_2nd_moment(z.to(torch.float32))
        for i in range(self.num_layers):
            x = getattr(self, "fc%d" % i)(x)
        if self.w_avg_beta is not None and self.training and not skip_w_avg_update:
            self.w_avg.copy_(x.detach().mean(dim=0).lerp(self.w_avg, self.w_avg_beta))
        if self.num_ws is not None:
            x = x.unsqueeze(1).repeat([1, self.num_ws, 1])
        if truncation_psi != 1:
            if self.num_ws is None or truncation_cutoff is None:
                x = self.w_avg.lerp(x, truncation_psi)
            else:
                x[:, :truncation_cutoff] = self.w_avg.lerp(x[:, :truncation_cutoff], truncation_psi)
        return x


def _up_img_pads(fk, up):
    """stylegan.py:620-632 / 769-781 with padding = 0: FIR pads around a x`up` or /`up` resample."""
    return [(fk + up - 1) // 2, (fk - up) // 2] * 3


class Conv3dLayer(tnn.Module):
    """stylegan.py:446-487: conv3d_resample with run-time weight gain, then bias + activation."""

    def __init__(self, in_channels, out_channels, kernel_size, bias=True, activation="linear", up=1, down=1,
                 resample_filter=(1, 3, 3, 1), trainable=True):
        super().__init__()
        self.activation, self.up, self.down = activation, up, down
        self.resample_filter = SO.make_resample_filter(resample_filter)
        self.padding = kernel_size // 2
        self.weight_gain = 1 / np.sqrt(in_channels * (kernel_size ** 3))
        weight = torch.randn([out_channels, in_channels, kernel_size, kernel_size, kernel_size])
        bias = torch.zeros([out_channels]) if bias else None
        if trainable:
            self.weight = tnn.Parameter(weight)
            self.bias = tnn.Parameter(bias) if bias is not None else None
        else:
            self.register_buffer("weight", weight)
            if bias is not None:
                self.register_buffer("bias", bias)
            else:
                self.bias = None

    def forward(self, x, gain=1):        # `gain` is ignored by the reference as well
        w = self.weight * self.weight_gain
        return SO.conv3d_resample(x=x, w=w, f=self.resample_filter, up=self.up, down=self.down,
                                  padding=self.padding, flip_weight=(self.up == 1), bias=self.bias,
                                  act=self.activation)


class SynthesisLayer(tnn.Module):
    """stylegan.py:489-532: style affine -> modulated conv (+ noise) -> bias -> lrelu."""

    def __init__(self, in_channels, out_channels, w_dim, resolution, kernel_size=3, up=1, use_noise=True,
                 activation="lrelu", resample_filter=(1, 3, 3, 1)):
        super().__init__()
        self.resolution, self.up, self.use_noise, self.activation = resolution, up, use_noise, activation
        self.resample_filter = SO.make_resample_filter(resample_filter)
        self.padding = kernel_size // 2
        self.affine = FullyConnectedLayer(w_dim, in_channels, bias_init=1)
        self.weight = tnn.Parameter(torch.randn([out_channels, in_channels, kernel_size, kernel_size, kernel_size]))
        if use_noise:
            self.register_buffer("noise_const", torch.randn([resolution, resolution]))
            self.noise_strength = tnn.Parameter(torch.zeros([]))
        self.bias = tnn.Parameter(torch.zeros([out_channels]))

    def forward(self, x, w, noise_mode="random", fused_modconv=False, gain=1, noise=None):
        """`noise` (extension, used by the parity tests): an explicit [N,1,R,R,R] N(0,1) field that
        replaces the `torch.randn` draw of noise_mode='random' (stylegan.py:522)."""
        styles = self.affine(w)
        field = None
        r = self.resolution
        if self.use_noise and noise_mode == "random":
            field = noise if noise is not None else torch.randn([x.shape[0], 1, r, r, r], device=x.device)
        if self.use_noise and noise_mode == "const":
            field = self.noise_const.expand(1, 1, r, r, r)          # [R, R] broadcasts over the two last axes
        return SO.synthesis_conv(x, self.weight, styles, field, self.noise_strength if field is not None else None,
                                 self.bias, up=self.up, padding=self.padding, resample_filter=self.resample_filter,
                                 act=self.activation)


class OutBlock(tnn.Module):
    """stylegan.py:534-546: 1x1x1 modulated conv to the image channels, no demodulation."""

    def __init__(self, in_channels, out_channels, w_dim, kernel_size=1):
        super().__init__()
        self.affine = FullyConnectedLayer(w_dim, in_channels, bias_init=1)
        self.weight = tnn.Parameter(torch.randn([out_channels, in_channels, kernel_size, kernel_size, kernel_size]))
        self.bias = tnn.Parameter(torch.zeros([out_channels]))
        self.weight_gain = 1 / np.sqrt(in_channels * (kernel_size ** 3))

    def forward(self, x, w, fused_modconv=False):
        styles = self.affine(w) * self.weight_gain
        if BF.mod_supported(x.shape[1], BF.compute_dtype(x)):
            return SO.conv3d_resample(BF.modulate(x, styles), self.weight, bias=self.bias)   # bias in the epilogue
        x = SO.modulated_conv3d(x=x, weight=self.weight, styles=styles, demodulate=False, fused_modconv=False)
        return SO.bias_act(x, self.bias)


class SynthesisBlock(tnn.Module):
    """stylegan.py:548-640 ('skip' architecture is what SynthesisNetwork builds)."""

    def __init__(self, in_channels, out_channels, w_dim, resolution, img_channels=1, is_last=False,
                 architecture="skip", resample_filter=(1, 3, 3, 1), use_fp16=False, fp16_channels_last=False,
                 **layer_kwargs):
        assert architecture in ("orig", "skip", "resnet")
        super().__init__()
        self.in_channels, self.w_dim, self.resolution = in_channels, w_dim, resolution
        self.img_channels, self.is_last, self.architecture = img_channels, is_last, architecture
        self.resample_filter = SO.make_resample_filter(resample_filter)
        self.num_conv = self.num_torgb = 0
        if in_channels == 0:
            self.const = tnn.Parameter(torch.randn([out_channels, resolution, resolution, resolution]))
        else:
            self.conv0 = SynthesisLayer(in_channels, out_channels, w_dim=w_dim, resolution=resolution, up=2,
                                        resample_filter=resample_filter, **layer_kwargs)
            self.num_conv += 1
        self.conv1 = SynthesisLayer(out_channels, out_channels, w_dim=w_dim, resolution=resolution, **layer_kwargs)
        self.num_conv += 1
        if is_last or architecture == "skip":
            self.torgb = OutBlock(out_channels, img_channels, w_dim=w_dim)
            self.num_torgb += 1
        if in_channels != 0 and architecture == "resnet":
            self.skip = Conv3dLayer(in_channels, out_channels, kernel_size=1, bias=False, up=2,
                                    resample_filter=resample_filter)

    def forward(self, x, ws, img=None, force_fp32=False, fused_modconv=None, noises=None, **layer_kwargs):
        w_iter = iter(ws.unbind(dim=1))
        nz = iter(noises) if noises is not None else None
        kw = lambda: dict(layer_kwargs, noise=next(nz)) if nz is not None else layer_kwargs   # noqa: E731
        if self.in_channels == 0:
            x = self.const.unsqueeze(0).repeat([ws.shape[0], 1, 1, 1, 1])
            x = self.conv1(x, next(w_iter), **kw())
        elif self.architecture == "resnet":
            y = self.skip(x)
            x = self.conv0(x, next(w_iter), **kw())
            x = self.conv1(x, next(w_iter), **kw())
            # (stylegan.py: `y.add_(x)`) out of place when y is a view - the W-folded skip conv returns one, and an
            # in-place op on a view makes autograd rebase it (CopySlices: a clone and two strided copies per pass)
            x = y + x if y._is_view() else y.add_(x)
        else:
            x = self.conv0(x, next(w_iter), **kw())
            x = self.conv1(x, next(w_iter), **kw())
        if img is not None:                      # x2 of the running image: zero-insert + FIR, gain 8
            fk = int(self.resample_filter.shape[0])
            img = BF.upfirdn3d(img, self.resample_filter, up=2, padding=_up_img_pads(fk, 2), gain=8)
        if self.is_last or self.architecture == "skip":
            y = self.torgb(x, next(w_iter)).to(dtype=torch.float32)
            img = img.add_(y) if img is not None else y
        return x, img


class SynthesisNetwork(tnn.Module):
    """stylegan.py:644-693."""

    def __init__(self, w_dim, img_resolution=128, img_channels=1, channel_base=4096, channel_max=512,
                 num_fp16_res=0, **block_kwargs):
        assert img_resolution >= 4 and img_resolution & (img_resolution - 1) == 0
        super().__init__()
        self.w_dim, self.img_resolution, self.img_channels = w_dim, img_resolution, img_channels
        self.img_resolution_log2 = int(np.log2(img_resolution))
        self.block_resolutions = [2 ** i for i in range(2, self.img_resolution_log2 + 1)]
        ch = {res: min(channel_base // res, channel_max) for res in self.block_resolutions}
        self.num_ws = 0
        for res in self.block_resolutions:
            block = SynthesisBlock(ch[res // 2] if res > 4 else 0, ch[res], w_dim=w_dim, resolution=res,
                                   img_channels=img_channels, is_last=(res == img_resolution), **block_kwargs)
            self.num_ws += block.num_conv
            if res == img_resolution:
                self.num_ws += block.num_torgb
            setattr(self, "b%d" % res, block)
        self.act = tnn.Tanh()

    def forward(self, ws, noises=None, **block_kwargs):
        """`noises` (extension for the parity tests): per block, the list of explicit noise fields."""
        ws = ws.to(torch.float32)
        x = img = None
        at = 0
        for i, res in enumerate(self.block_resolutions):
            block = getattr(self, "b%d" % res)
            cur = ws.narrow(1, at, block.num_conv + block.num_torgb)
            at += block.num_conv
            x, img = block(x, cur, img, noises=None if noises is None else noises[i], **block_kwargs)
        return BF.activation(img, "tanh")


class Generator(tnn.Module):
    """stylegan.py:697-718: returns (img, ws)."""

    def __init__(self, params, w_dim=512, img_resolution=128, img_channels=1, mapping_kwargs={}, synthesis_kwargs={}):
        super().__init__()
        self.p = params
        self.z_dim, self.w_dim = params.z_size, w_dim
        self.img_resolution, self.img_channels = img_resolution, img_channels
        self.synthesis = SynthesisNetwork(w_dim=w_dim, img_resolution=img_resolution, channel_base=params.filterG,
                                          **synthesis_kwargs)
        self.num_ws = self.synthesis.num_ws
        self.mapping = MappingNetwork(z_dim=self.z_dim, w_dim=w_dim, num_ws=self.num_ws, **mapping_kwargs)

    def forward(self, z, **synthesis_kwargs):
        ws = self.mapping(z)
        return self.synthesis(ws, **synthesis_kwargs), ws


class DiscriminatorBlock(tnn.Module):
    """stylegan.py:721-811 ('resnet' architecture is the default the Discriminator builds)."""

    def __init__(self, in_channels, tmp_channels, out_channels, resolution, img_channels, first_layer_idx,
                 architecture="resnet", activation="lrelu", resample_filter=(1, 3, 3, 1), conv_clamp=None,
                 use_fp16=False, fp16_channels_last=False, freeze_layers=0):
        super().__init__()
        self.in_channels, self.resolution, self.img_channels = in_channels, resolution, img_channels
        self.first_layer_idx, self.architecture = first_layer_idx, architecture
        self.resample_filter = SO.make_resample_filter(resample_filter)
        self.num_layers = 0

        def trainable():
            t = (self.first_layer_idx + self.num_layers) >= freeze_layers
            self.num_layers += 1
            return t

        if in_channels == 0 or architecture == "skip":
            self.fromrgb = Conv3dLayer(img_channels, tmp_channels, kernel_size=1, activation=activation,
                                       trainable=trainable())
        self.conv0 = Conv3dLayer(tmp_channels, tmp_channels, kernel_size=3, activation=activation,
                                 trainable=trainable())
        self.conv1 = Conv3dLayer(tmp_channels, out_channels, kernel_size=3, activation=activation, down=2,
                                 trainable=trainable(), resample_filter=resample_filter)
        if architecture == "resnet":
            self.skip = Conv3dLayer(tmp_channels, out_channels, kernel_size=1, bias=False, down=2,
                                    trainable=trainable(), resample_filter=resample_filter)

    def forward(self, x, img, force_fp32=False):
        if self.in_channels == 0 or self.architecture == "skip":
            y = self.fromrgb(img)
            x = x + y if x is not None else y
            if self.architecture == "skip":
                fk = int(self.resample_filter.shape[0])
                img = BF.upfirdn3d(x, self.resample_filter, down=2,
                                   padding=[(fk - 2 + 1) // 2, (fk - 2) // 2] * 3, flip_filter=False)
            else:
                img = None
        if self.architecture == "resnet":
            y = self.skip(x)
            x = self.conv0(x)
            x = self.conv1(x)
            # (stylegan.py: `y.add_(x)`) out of place when y is a view - the W-folded skip conv returns one, and an
            # in-place op on a view makes autograd rebase it (CopySlices: a clone and two strided copies per pass)
            x = y + x if y._is_view() else y.add_(x)
        else:
            x = self.conv0(x)
            x = self.conv1(x)
        return x, img


class MinibatchStdLayer(tnn.Module):
    """stylegan.py:814-835."""

    def __init__(self, group_size, num_channels=1):
        super().__init__()
        self.group_size, self.num_channels = group_size, num_channels

    def forward(self, x):
        n, c, a0, a1, a2 = x.shape
        g = min(int(self.group_size), n) if self.group_size is not None else n
        f = self.num_channels
        y = x.float().reshape(g, -1, f, c // f, a0, a1, a2)
        y = y - y.mean(dim=0)
        y = (y.square().mean(dim=0) + 1e-8).sqrt()
        y = y.mean(dim=[2, 3, 4, 5]).reshape(-1, f, 1, 1, 1)
        y = y.repeat(g, 1, a0, a1, a2).to(x.dtype)
        return torch.cat([x, y], dim=1)


class DiscriminatorEpilogue(tnn.Module):
    """stylegan.py:838-880."""

    def __init__(self, in_channels, resolution, img_channels, architecture="resnet", mbstd_group_size=4,
                 mbstd_num_channels=1, activation="lrelu", conv_clamp=None):
        assert architecture in ("orig", "skip", "resnet")
        super().__init__()
        self.in_channels, self.resolution = in_channels, resolution
        self.img_channels, self.architecture = img_channels, architecture
        if architecture == "skip":
            self.fromrgb = Conv3dLayer(img_channels, in_channels, kernel_size=1, activation=activation)
        self.mbstd = MinibatchStdLayer(mbstd_group_size, mbstd_num_channels) if mbstd_num_channels > 0 else None
        self.conv = Conv3dLayer(in_channels + mbstd_num_channels, in_channels, kernel_size=3, activation=activation)
        self.fc = FullyConnectedLayer(in_channels * (resolution ** 3), in_channels, activation=activation)
        self.out = FullyConnectedLayer(in_channels, 1)

    def _conv_padded(self, x):
        """`in_channels + 1` (513) input channels have no 16-byte channel vectors: zero-pad the
        activation and the filter to the next multiple of 8 channels so the layer takes the
        tcgen05 path (the padded taps multiply zeros; their gradient is sliced away by autograd)."""
        extra = (-x.shape[1]) % 8
        if extra == 0 or x.device.type != "cuda":
            return self.conv(x)
        c = self.conv
        x = torch.cat([x, x.new_zeros(x.shape[0], extra, *x.shape[2:])], dim=1)
        w = F.pad(c.weight * c.weight_gain, (0, 0, 0, 0, 0, 0, 0, extra))
        return SO.conv3d_resample(x=x, w=w, f=c.resample_filter, up=c.up, down=c.down, padding=c.padding,
                                  flip_weight=(c.up == 1), bias=c.bias, act=c.activation)

    def forward(self, x, img, force_fp32=False):
        if self.architecture == "skip":
            x = x + self.fromrgb(img)
        if self.mbstd is not None:
            x = self.mbstd(x)
        x = self._conv_padded(x)
        x = self.fc(x.flatten(1).float())
        return self.out(x)


class Discriminator(tnn.Module):
    """stylegan.py:883-928."""

    def __init__(self, params, img_resolution=128, img_channels=1, architecture="resnet", channel_max=512,
                 num_fp16_res=0, conv_clamp=None, cmap_dim=None, block_kwargs={}, mapping_kwargs={},
                 epilogue_kwargs={}):
        super().__init__()
        channel_base = params.filterD
        self.img_resolution, self.img_channels = img_resolution, img_channels
        self.img_resolution_log2 = int(np.log2(img_resolution))
        self.block_resolutions = [2 ** i for i in range(self.img_resolution_log2, 2, -1)]
        ch = {res: min(channel_base // res, channel_max) for res in self.block_resolutions + [4]}
        common = dict(img_channels=img_channels, architecture=architecture)
        idx = 0
        for res in self.block_resolutions:
            block = DiscriminatorBlock(ch[res] if res < img_resolution else 0, ch[res], ch[res // 2], resolution=res,
                                       first_layer_idx=idx, **block_kwargs, **common)
            setattr(self, "b%d" % res, block)
            idx += block.num_layers
        self.b4 = DiscriminatorEpilogue(ch[4], resolution=4, **epilogue_kwargs, **common)

    def forward(self, img, **block_kwargs):
        x = None
        for res in self.block_resolutions:
            x, img = getattr(self, "b%d" % res)(x, img, **block_kwargs)
        return self.b4(x, img)


# ------------------------------------------------------------------ StyleGAN-v1-like generator
def instance_stats(f, eps=1e-5):
    """stylegan.py:931-941: per (sample, channel) mean and sqrt(unbiased var + eps)."""
    n, c = f.shape[:2]
    flat = f.float().reshape(n, c, -1)
    var, mean = torch.var_mean(flat, dim=2)
    return mean.view(n, c, 1, 1, 1), (var + eps).sqrt().view(n, c, 1, 1, 1)


def AdaIN(f_content, f_style, up=False):
    """stylegan.py:943-959: instance-normalise, then scale / shift by the style's (mean | std) halves.
    `up=True` folds the `F.interpolate(scale_factor=2, mode='trilinear')` that follows every second
    AdaIN (stylegan.py:1081-1137) into the same pass (interpolation commutes with the per-(n,c) affine)."""
    nc = f_content.shape[1]
    if f_content.is_cuda and BF.mod_supported(nc, BF.compute_dtype(f_content)):
        return BF.adain(f_content, f_style[:, 0:nc], f_style[:, nc:], 1e-5, up)
    mean, std = instance_stats(f_content)
    s_mean = f_style[:, 0:nc].float()[:, :, None, None, None]
    s_std = f_style[:, nc:].float()[:, :, None, None, None]
    a = s_std / std
    b = s_mean - mean * a
    y = torch.addcmul(b.to(f_content.dtype), f_content, a.to(f_content.dtype))
    return F.interpolate(y, scale_factor=2, mode="trilinear") if up else y


def conditionalSplit(w, swapPoint, layerCtr, alreadySplit):
    """stylegan.py:961-967: batch-shuffle style mixing."""
    if (layerCtr == swapPoint) and not alreadySplit:
        w = w[torch.randperm(w.shape[0])]
    return w


class styleGAN_gen(tnn.Module):
    """stylegan.py:969-1149: 8-layer nn.Linear mapping, constant ones [B,512,4,4,4], ten
    `Conv3d(k3,p1,no bias) + LeakyReLU(0.2)` with AdaIN between, trilinear x2 five times,
    `Conv3d(16->1) + Tanh`.  Convolutions are the B200 drop-in layers with the activation fused
    into the epilogue."""

    WIDTHS = [(512, 512), (512, 256), (256, 256), (256, 128), (128, 128), (128, 64), (64, 64), (64, 32),
              (32, 32), (32, 16)]

    def __init__(self, params, styleMixing=True, mappingNet=True):
        super().__init__()
        self.p = params
        self.nz = params.z_size
        self.styleMixing, self.mappingNet = styleMixing, mappingNet
        if mappingNet:
            layers = []
            for _ in range(8):
                layers += [tnn.Linear(self.nz, self.nz), tnn.LeakyReLU(0.2, inplace=True)]
            self.latentMapping = tnn.Sequential(*layers)
        for i, (ci, co) in enumerate(self.WIDTHS):
            conv = BN.Conv3d(ci, co, (3, 3, 3), (1, 1, 1), (1, 1, 1), bias=False)
            setattr(self, "C%d" % (i + 1), tnn.Sequential(conv, BN.LeakyReLU(0.2, inplace=True)))
        self.C_out = tnn.Sequential(BN.Conv3d(16, 1, (3, 3, 3), (1, 1, 1), (1, 1, 1), bias=False), BN.Tanh())
        for i, c in enumerate([512, 512, 256, 256, 128, 128, 64, 64, 32, 32, 16]):
            setattr(self, "A%d" % (i + 1), tnn.Linear(512, 2 * c))
        from .patch import fuse
        fuse(self)

    def forward(self, z_in, w_passed=False):
        n, c = z_in.shape[0], z_in.shape[1]
        assert c == 512
        mixing = self.styleMixing and self.training
        swap = torch.randint(6, (1, 1, 1)).item() if mixing else None
        ctr = 0
        if (not self.mappingNet) or w_passed:
            w = torch.squeeze(z_in)
        else:
            w = self.latentMapping(torch.squeeze(z_in))
        h = torch.ones((n, c, 4, 4, 4), dtype=z_in.dtype, device=z_in.device)
        h = AdaIN(h, self.A1(w))
        for i in range(1, 11):
            h = getattr(self, "C%d" % i)(h)
            if mixing and (i % 2 == 1 or i == 10):     # after C1, C3, C5, C7, C9 and C10: style shuffle point
                w = conditionalSplit(w, swap, ctr, False)
                ctr += 1
            h = AdaIN(h, getattr(self, "A%d" % (i + 1))(w), up=(i % 2 == 1))
        return self.C_out(h)


# ------------------------------------------------------------------ losses (stylegan.py:6-100)
class StyleGAN2Loss:
    """stylegan.py:6-100: non-saturating logistic losses, R1 on every 16th step (every step for the
    v1-like generator), path-length regularisation on every 16th generator step (StyleGAN2 only).
    `G_synthesis` is None for the v1-like generator, whose forward returns the image only."""

    def __init__(self, params, G_mapping, G_synthesis, D, augment_pipe=None, style_mixing_prob=0.9, r1_gamma=10,
                 pl_batch_shrink=2, pl_decay=0.01, pl_weight=2):
        self.p, self.device = params, params.device
        self.G_mapping, self.G_synthesis, self.D = G_mapping, G_synthesis, D
        self.style_mixing_prob, self.r1_gamma = style_mixing_prob, r1_gamma
        self.pl_batch_shrink, self.pl_decay, self.pl_weight = pl_batch_shrink, pl_decay, pl_weight
        self.pl_mean = torch.zeros([], device=self.device)

    def run_G(self, z):
        if self.p.stylegan2:
            ws = self.G_mapping(z)
            if self.style_mixing_prob > 0:
                cutoff = torch.empty([], dtype=torch.int64, device=ws.device).random_(1, ws.shape[1])
                cutoff = torch.where(torch.rand([], device=ws.device) < self.style_mixing_prob, cutoff,
                                     torch.full_like(cutoff, ws.shape[1]))
                ws[:, cutoff:] = self.G_mapping(torch.randn_like(z), skip_w_avg_update=True)[:, cutoff:]
            return self.G_synthesis(ws), ws
        return self.G_mapping(z)

    def run_D(self, img):
        return self.D(img)

    def step_D(self, step, real_img, gen_z, gain=1):
        do_r1 = ((step % 16) == 0 and self.r1_gamma != 0) or (self.p.stylegan and self.r1_gamma != 0)
        gen_img, _ = self.run_G(gen_z)
        loss_gen = F.softplus(self.run_D(gen_img).float())
        loss_gen.mean().mul(gain).backward()
        real = real_img.detach().requires_grad_(do_r1)
        real_logits = self.run_D(real).float()
        loss_real = F.softplus(-real_logits)
        loss_r1 = 0
        if do_r1:
            (g,) = torch.autograd.grad(outputs=[real_logits.sum()], inputs=[real], create_graph=False,
                                       retain_graph=True, only_inputs=True)
            loss_r1 = g.float().square().sum([1, 2, 3]) * (self.r1_gamma / 2)
        total = (real_logits * 0 + loss_real + loss_r1).mean().mul(gain)
        total.backward()
        return total.detach(), loss_gen.mean().mul(gain).detach()

    def step_G(self, step, gen_z, gain=1, pl_noise=None):
        """`pl_noise` (extension for the parity tests) replaces the randn_like draw of stylegan.py:86."""
        do_pl = (step % 16) == 0 and self.p.stylegan2
        gen_img, _ = self.run_G(gen_z)
        loss_main = F.softplus(-self.run_D(gen_img).float())
        loss_main.mean().mul(gain).backward()
        loss = loss_main.mean().mul(gain).detach()
        if do_pl:
            n = gen_z.shape[0] // self.pl_batch_shrink
            gen_img, gen_ws = self.run_G(gen_z[:n])
            if pl_noise is None:
                pl_noise = torch.randn_like(gen_img)
            pl_noise = pl_noise.to(gen_img.dtype) / math.sqrt(gen_img.shape[2] * gen_img.shape[3])
            (g,) = torch.autograd.grad(outputs=[(gen_img * pl_noise).sum()], inputs=[gen_ws], create_graph=False,
                                       retain_graph=True, only_inputs=True)
            lengths = g.float().square().sum(2).mean(1).sqrt()
            pl_mean = self.pl_mean.lerp(lengths.mean(), self.pl_decay)
            self.pl_mean.copy_(pl_mean.detach())
            loss_pl = ((lengths - pl_mean).square() * self.pl_weight)
            total = (gen_img[:, 0, 0, 0, 0] * 0 + loss_pl).mean().mul(gain)
            total.backward()
            loss = total.detach() + loss
        return loss, gen_img
