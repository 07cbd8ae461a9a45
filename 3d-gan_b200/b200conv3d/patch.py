"""The single switch that turns the replacement on (SURVEY.md §5 "Config").

* :func:`install` rebinds ``torch.nn.Conv3d / ConvTranspose3d / BatchNorm3d /
  LayerNorm / ReLU / LeakyReLU / Tanh`` and ``torch.nn.functional.conv3d /
  conv_transpose3d`` so that the reference's model files (dcgan.py, biggan.py,
  utils.py, stylegan.py, FID_ResNet.py), which only say ``nn.Conv3d(...)``,
  build B200 layers without being edited.
* :func:`convert` retrofits an already-built model in place (class swap; the
  parameters are untouched so ``state_dict`` round-trips both ways).
* :func:`fuse` folds ``X -> activation`` pairs inside ``nn.Sequential``s into
  X's epilogue without renaming anything.
"""
import torch
import torch.nn as tnn
import torch.nn.functional as TF

from . import functional as BF
from . import nn as BN

_CLASS_MAP = {
    "Conv3d": (BN._TConv3d, BN.Conv3d),
    "ConvTranspose3d": (BN._TConvTranspose3d, BN.ConvTranspose3d),
    "BatchNorm3d": (BN._TBatchNorm3d, BN.BatchNorm3d),
    "LayerNorm": (BN._TLayerNorm, BN.LayerNorm),
    "ReLU": (BN._TReLU, BN.ReLU),
    "LeakyReLU": (BN._TLeakyReLU, BN.LeakyReLU),
    "Tanh": (BN._TTanh, BN.Tanh),
    "MaxPool3d": (BN._TMaxPool3d, BN.MaxPool3d),          # FID_ResNet.py:129
    "AvgPool3d": (BN._TAvgPool3d, BN.AvgPool3d),          # biggan.py:86,92
}
_stock_functional = {"conv3d": TF.conv3d, "conv_transpose3d": TF.conv_transpose3d, "interpolate": TF.interpolate}
_installed = [False]


def _f_conv3d(input, weight, bias=None, stride=1, padding=0, dilation=1, groups=1):
    return BF.conv3d(input, weight, bias, stride, padding, dilation, groups)


def _f_conv_transpose3d(input, weight, bias=None, stride=1, padding=0, output_padding=0,
                        groups=1, dilation=1):
    return BF.conv_transpose3d(input, weight, bias, stride, padding, output_padding, groups,
                               dilation)


def _f_interpolate(input, size=None, scale_factor=None, mode="nearest", align_corners=None,
                   recompute_scale_factor=None, antialias=False):
    """``F.interpolate(x, scale_factor=2)`` on a 5-D tensor (nearest; biggan.py:26,31 through utils.py:78-80) is one
    NDHWC pass of the library.  Every other use of F.interpolate keeps torch's implementation: it is not a
    convolution and not on the path this package replaces (trilinear x2 of stylegan.py:1081 is SURVEY 8(f) n1)."""
    sf = scale_factor
    if isinstance(sf, (list, tuple)) and len(set(sf)) == 1:
        sf = sf[0]
    if (input.dim() == 5 and mode == "nearest" and size is None and sf is not None and not isinstance(sf, (list, tuple))
            and float(sf) == 2.0 and (input.is_cuda or BF._lib.is_host())):
        return BF.upsample_nearest2x(input)
    return _stock_functional["interpolate"](input, size=size, scale_factor=scale_factor, mode=mode,
                                            align_corners=align_corners,
                                            recompute_scale_factor=recompute_scale_factor, antialias=antialias)


def install():
    """Rebind the torch layer names the reference uses.  Idempotent."""
    for name, (_, ours) in _CLASS_MAP.items():
        setattr(tnn, name, ours)
    TF.conv3d = _f_conv3d
    TF.conv_transpose3d = _f_conv_transpose3d
    TF.interpolate = _f_interpolate
    BF.watch_optimizers()
    BF.invalidate_packs()
    _installed[0] = True


def uninstall():
    for name, (stock, _) in _CLASS_MAP.items():
        setattr(tnn, name, stock)
    TF.conv3d = _stock_functional["conv3d"]
    TF.conv_transpose3d = _stock_functional["conv_transpose3d"]
    TF.interpolate = _stock_functional["interpolate"]
    _installed[0] = False


def is_installed():
    return _installed[0]


def _swap_class(m):
    cls = type(m)
    for stock, ours in _CLASS_MAP.values():
        if cls is stock:
            m.__class__ = ours
            return True
        # spectral_norm & co. give each module its own "Parametrized<Cls>" class
        if cls.__name__.startswith("Parametrized") and cls.__bases__ == (stock,):
            cls.__bases__ = (ours,)
            return True
    return False


def convert(model):
    """Swap stock torch layers inside ``model`` for B200 layers, in place."""
    n = 0
    for m in model.modules():
        n += bool(_swap_class(m))
    BF.invalidate_packs()          # parameters may have been loaded / broadcast through `.data` before the swap
    return n


def _act_of(m):
    if isinstance(m, BN._TLeakyReLU):
        return "lrelu", float(m.negative_slope)
    if isinstance(m, BN._TReLU):
        return "relu", 0.0
    if isinstance(m, BN._TTanh):
        return "tanh", 0.0
    return None


def _fusable(m):
    if isinstance(m, (BN.Conv3d, BN.ConvTranspose3d, BN.BatchNorm3d)):
        return m.fused_act is None
    if isinstance(m, BN.LayerNorm):
        return m.fused_act is None and len(m.normalized_shape) == 4
    return False


def fuse(model):
    """Fold ``layer -> ReLU/LeakyReLU/Tanh`` pairs of every nn.Sequential into
    the layer's epilogue (dcgan.py:45-70: ConvT-BN-ReLU, ConvT-Tanh;
    dcgan.py:117-182: Conv-LN-LeakyReLU, Conv-LeakyReLU).  Returns the number
    of activations folded.  Indices, names and parameters do not change."""
    n = 0
    for seq in model.modules():
        if not isinstance(seq, tnn.Sequential):
            continue
        for i in range(len(seq) - 1):
            a, b = seq[i], seq[i + 1]
            act = _act_of(b)
            if act is not None and _fusable(a):
                a.fused_act, a.fused_slope = act
                seq[i + 1] = BN.FusedAway(b)
                n += 1
            # conv -> BatchNorm3d / whole-sample LayerNorm: the statistics ride in the conv's GEMM epilogue
            if isinstance(a, (BN.Conv3d, BN.ConvTranspose3d)) and a.fused_act is None:
                if isinstance(b, BN.BatchNorm3d):
                    a.fused_stats = "channel"
                elif isinstance(b, BN.LayerNorm) and len(b.normalized_shape) == 4:
                    a.fused_stats = "sample"
    return n


def prepare(model, do_fuse=True):
    """convert + fuse; returns the model."""
    BF.watch_optimizers()
    convert(model)
    if do_fuse:
        fuse(model)
    return model
