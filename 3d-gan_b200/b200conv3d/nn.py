"""Drop-in ``torch.nn`` layer subclasses (Face 1).

Each class subclasses the torch layer it replaces, so constructor signature,
parameter / buffer names, shapes, dtypes and ``state_dict`` are identical, and
``isinstance(m, nn.Conv3d)`` (biggan.py:49,105; FID_ResNet.py:140-142),
``m.__class__.__name__.find('Conv')`` (dcgan.py:75-81) and
``spectral_norm(module)`` (utils.py:13) keep working.  Only ``forward`` differs:
it calls the sm_100a library through :mod:`b200conv3d.functional`.

``fused_act`` / ``fused_slope`` are set by :func:`b200conv3d.patch.fuse` when the
next module of an ``nn.Sequential`` is an activation; they are plain Python
attributes, never parameters.
"""
import torch
import torch.nn as tnn

from . import functional as BF

# the stock classes, captured before install() can rebind torch.nn.*
_TConv3d = tnn.Conv3d
_TConvTranspose3d = tnn.ConvTranspose3d
_TBatchNorm3d = tnn.BatchNorm3d
_TLayerNorm = tnn.LayerNorm
_TReLU = tnn.ReLU
_TLeakyReLU = tnn.LeakyReLU
_TTanh = tnn.Tanh


class Conv3d(_TConv3d):
    """nn.Conv3d replacement (dcgan.py:161-181, utils.py:8-14,
    FID_ResNet.py:9-16,70-75,119-125,157-162, stylegan.py:998-1038)."""
    fused_act = None
    fused_slope = 0.0
    fused_stats = None        # "channel" / "sample": the next layer is a BatchNorm3d / whole-sample LayerNorm

    def forward(self, x):
        if self.padding_mode != "zeros":
            raise NotImplementedError("b200conv3d.Conv3d: padding_mode=%r is not supported"
                                      % self.padding_mode)
        return BF.conv3d(x, self.weight, self.bias, self.stride, self.padding, self.dilation,
                         self.groups, act=self.fused_act, slope=self.fused_slope, stats=self.fused_stats)


class ConvTranspose3d(_TConvTranspose3d):
    """nn.ConvTranspose3d replacement (dcgan.py:19-40,47-67)."""
    fused_act = None
    fused_slope = 0.0
    fused_stats = None

    def forward(self, x, output_size=None):
        if self.padding_mode != "zeros":
            raise NotImplementedError("b200conv3d.ConvTranspose3d: only zeros padding")
        out_pad = self.output_padding
        if output_size is not None:
            out_pad = self._output_padding(x, output_size, self.stride, self.padding,
                                           self.kernel_size, 3, self.dilation)
        return BF.conv_transpose3d(x, self.weight, self.bias, self.stride, self.padding, out_pad,
                                   self.groups, self.dilation, act=self.fused_act,
                                   slope=self.fused_slope, stats=self.fused_stats)


class BatchNorm3d(_TBatchNorm3d):
    """nn.BatchNorm3d replacement (dcgan.py:48-64, utils.py:62-65, biggan.py:40,
    FID_ResNet.py:71-76,127,162).  Same momentum / running-stat semantics as
    torch (including ``num_batches_tracked`` and ``momentum=None``)."""
    fused_act = None
    fused_slope = 0.0

    def forward(self, x, residual=None):
        self._check_input_dim(x)
        eaf = 0.0 if self.momentum is None else self.momentum
        if self.training and self.track_running_stats and self.num_batches_tracked is not None:
            self.num_batches_tracked.add_(1)
            if self.momentum is None:
                eaf = 1.0 / float(self.num_batches_tracked)
        bn_training = self.training or (self.running_mean is None and self.running_var is None)
        rm = self.running_mean if (not self.training or self.track_running_stats) else None
        rv = self.running_var if (not self.training or self.track_running_stats) else None
        return BF.batch_norm_act(x, self.weight, self.bias, rm, rv, bn_training, eaf, self.eps,
                                 act=self.fused_act, slope=self.fused_slope, residual=residual)


class LayerNorm(_TLayerNorm):
    """nn.LayerNorm replacement for the whole-sample form the DCGAN
    discriminator uses, ``LayerNorm([C,D,H,W])`` on a 5-D input
    (dcgan.py:162,166,170,174,178).  Any other use of nn.LayerNorm is not on
    the reference's hot path and keeps torch's implementation."""
    fused_act = None
    fused_slope = 0.0

    def forward(self, x):
        if x.dim() == 5 and tuple(self.normalized_shape) == tuple(x.shape[1:]):
            return BF.layer_norm_act(x, self.weight, self.bias, self.eps, act=self.fused_act,
                                     slope=self.fused_slope)
        if self.fused_act is not None:
            raise RuntimeError("fused LayerNorm activation needs the whole-sample 5-D form")
        return super().forward(x)


class ReLU(_TReLU):
    def forward(self, x):
        return BF.activation(x, "relu", 0.0, self.inplace)


class LeakyReLU(_TLeakyReLU):
    def forward(self, x):
        return BF.activation(x, "lrelu", self.negative_slope, self.inplace)


class Tanh(_TTanh):
    def forward(self, x):
        return BF.activation(x, "tanh", 0.0, False)


_TAvgPool3d = tnn.AvgPool3d


class AvgPool3d(_TAvgPool3d):
    """nn.AvgPool3d(2) of the BigGAN discriminator blocks (utils.py:113-116, biggan.py:86,92) as
    one NDHWC pass.  Other pooling configurations are not on the reference's path: they raise."""

    def forward(self, x):
        k, s, p = BF._triple(self.kernel_size), BF._triple(self.stride), BF._triple(self.padding)
        if k != (2, 2, 2) or s != (2, 2, 2) or p != (0, 0, 0) or self.ceil_mode or self.divisor_override:
            raise ValueError("b200conv3d.nn.AvgPool3d implements kernel_size=2, stride=2, padding=0 only")
        return BF.avg_pool2(x)


_TMaxPool3d = tnn.MaxPool3d


class MaxPool3d(_TMaxPool3d):
    """nn.MaxPool3d(kernel_size=3, stride=2, padding=1) after the FID stem (FID_ResNet.py:129) as one NDHWC pass."""

    def forward(self, x):
        if self.ceil_mode or self.return_indices or BF._triple(self.dilation) != (1, 1, 1):
            raise ValueError("b200conv3d.nn.MaxPool3d: ceil_mode / return_indices / dilation are not on the reference's path")
        return BF.max_pool3d(x, self.kernel_size, self.stride, self.padding)


class FusedAway(tnn.Identity):
    """Placeholder left where an activation was folded into the previous layer
    by :func:`b200conv3d.patch.fuse`; keeps nn.Sequential indices (and thus
    ``state_dict`` keys such as ``main.3.weight``) unchanged."""

    def __init__(self, original):
        super().__init__()
        self.original_repr = repr(original)

    def extra_repr(self):
        return "fused: " + self.original_repr
