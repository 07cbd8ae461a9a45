"""The StyleGAN2 convolution funnel of the reference on the B200 path (SURVEY.md §8(a) rows
a9-a12): every convolution of `stylegan.py` goes through `conv3d_resample`
(stylegan.py:202-294) and, in the generator, `modulated_conv3d` (stylegan.py:396-444).

Same names, argument meaning and results as the reference for the arguments its layers pass
(isotropic int `up` / `down` / `padding`, 3-D resample filter, `groups=1`,
`fused_modconv=False` - the grouped "fused" branch is dead code in training and crashes with
demodulation).  What runs where:

  * dense convolutions / transposed convolutions -> b2c_conv3d_* (tcgen05 path in bf16);
  * pad + depthwise 4x4x4 FIR + decimation / zero-insertion (`_upfirdn3d_ref`,
    stylegan.py:172-200) -> ONE b2c_fir3d pass (no materialised padding, no groups=C conv);
  * style scaling of the input -> ONE b2c_modulate pass; demodulation, noise, bias and activation ->
    ONE b2c_mod_epilogue pass (`synthesis_conv`); the demodulation coefficients are a small fp32 GEMM;
  * narrow stride-2 up-sampling convs (and the data gradient of the down-sampling convs) -> the
    phase-merged form of functional.conv_transpose3d_s2: one stride-1 `Cin -> 8*Cout` conv +
    b2c_depth_to_space2;  `fromrgb` / `torgb` 1x1x1 convs -> the pointwise kernels of csrc/thin.cu.

With `b200conv3d.install()` the reference's own stylegan.py reaches the same kernels through
`F.conv3d` / `F.conv_transpose3d`, except its `groups=C` FIR call, which this module replaces.
"""
import torch

from . import functional as BF


def make_resample_filter(taps=(1, 3, 3, 1)):
    """The reference's (non-separable, "odd") 4x4x4 low-pass filter, stylegan.py:462-465."""
    f = torch.tensor([float(t) for t in taps])
    f = (f[None, None, :] * f[None, :, None]).repeat((1, 4, 1)).reshape(4, 4, 4)
    return f / f.sum()


_ACT = {"linear": (None, 0.0), "lrelu": ("lrelu", 0.2), "relu": ("relu", 0.0), "tanh": ("tanh", 0.0)}


def _conv(x, w, stride=1, padding=0, transpose=False, flip_weight=True, bias=None, act="linear"):
    """stylegan.py:164-170.  `flip_weight=False` means true convolution: reverse the taps.
    `bias` / `act` ride in the GEMM epilogue (bias_act of stylegan.py:118-124 fused)."""
    if not flip_weight:
        w = w.flip([2, 3, 4])
    if act not in _ACT:
        raise NotImplementedError("activation %r is not used by the reference's networks" % (act,))
    a, slope = _ACT[act]
    if transpose:
        return BF.conv_transpose3d(x, w, bias, stride, padding, act=a, slope=slope)
    return BF.conv3d(x, w, bias, stride, padding, act=a, slope=slope)


def _resample_pads(fk, up, down, padding):
    """Padding (before, after) relative to the up-sampled grid once the FIR taps are accounted
    for (stylegan.py:228-241)."""
    lo = hi = int(padding)
    if up > 1:
        lo, hi = lo + (fk + up - 1) // 2, hi + (fk - up) // 2
    if down > 1:
        lo, hi = lo + (fk - down + 1) // 2, hi + (fk - down) // 2
    return lo, hi


def conv3d_resample(x, w, f=None, up=1, down=1, padding=0, groups=1, flip_weight=True, flip_filter=False,
                    bias=None, act="linear"):
    """stylegan.py:202-294.  Extension: `bias` / `act` = the `bias_act` the calling layer applies to
    the result (stylegan.py:485); fused into the convolution's epilogue whenever the convolution is
    the last operation of the funnel, applied as a separate pass after the FIR otherwise."""
    if groups != 1:
        raise NotImplementedError("grouped conv3d_resample is only reached by the dead fused_modconv branch")
    if not all(isinstance(v, int) for v in (up, down, padding)):
        raise NotImplementedError("the reference passes one int for up / down / padding")
    k = int(w.shape[2])
    fk = 1 if f is None else int(f.shape[0])
    lo, hi = _resample_pads(fk, up, down, padding)
    fir = lambda t, **kw: BF.upfirdn3d(t, f, flip_filter=flip_filter, **kw)   # noqa: E731
    pads = lambda a, b: [a, b, a, b, a, b]                                    # noqa: E731

    ep = dict(bias=bias, act=act)
    if up == 1 and down == 1:
        if lo == hi and lo >= 0:
            return _conv(x, w, padding=lo, flip_weight=flip_weight, **ep)
        return _conv(BF.upfirdn3d(x, None, padding=pads(lo, hi)), w, flip_weight=flip_weight, **ep)
    if up == 1:                                  # down-sampling
        if k == 1:                               # decimate first, then the 1x1x1 conv (:244-247)
            return _conv(fir(x, down=down, padding=pads(lo, hi)), w, flip_weight=flip_weight, **ep)
        return _conv(fir(x, padding=pads(lo, hi)), w, stride=down, flip_weight=flip_weight, **ep)   # (:257-260)
    if k == 1 and down == 1:                     # 1x1x1 conv first, then zero-insert + FIR (:250-254)
        y = fir(_conv(x, w, flip_weight=flip_weight), up=up, padding=pads(lo, hi), gain=up ** 2)
        return bias_act(y, bias, act=act)
    # up-sampling conv = transposed conv with stride `up`, then the FIR on what is left of the
    # padding (:262-282); the reference's FIR gain is up**2 in 3-D as well - kept
    lo, hi = lo - (k - 1), hi - (k - up)
    trim = max(min(-lo, -hi), 0)
    wt = w.transpose(0, 1)
    if up == 2 and trim == 0 and hi >= 1 and BF.phase_merge_applies(x, wt):
        # narrow up-sampling conv: the 8 sub-pixel phases as ONE stride-1 conv + depth-to-space, on the even
        # grid whose last plane is zero - the FIR takes one voxel less of trailing padding instead
        y = BF.conv_transpose3d_s2(x, wt.flip([2, 3, 4]) if flip_weight else wt, even=True)
        y = BF.upfirdn3d(y, f, flip_filter=flip_filter, padding=[lo, hi - 1] * 3, gain=up ** 2)
    else:
        y = _conv(x, wt, stride=up, padding=trim, transpose=True, flip_weight=not flip_weight)
        y = fir(y, padding=pads(lo + trim, hi + trim), gain=up ** 2)
    if down > 1:
        y = fir(y, down=down)
    return bias_act(y, bias, act=act)


def demod_coefficients(weight, styles):
    """dcoefs[n, o] = rsqrt(sum_{i,k} (w[o,i,k] * s[n,i])^2 + 1e-8)  (stylegan.py:419-422), as the
    small GEMM  s^2 [N, I] x (sum_k w^2)^T [I, O]  instead of an [N, O, I, k^3] temporary."""
    w2 = weight.float().square().sum(dim=[2, 3, 4])            # [O, I]
    with torch.autocast(styles.device.type, enabled=False):    # a [N,I] x [I,O] GEMM: keep it out of bf16 autocast
        return (styles.float().square() @ w2.t() + 1e-8).rsqrt()   # [N, O]


def modulated_conv3d(x, weight, styles, noise=None, up=1, down=1, padding=0, resample_filter=None,
                     demodulate=True, flip_weight=True, fused_modconv=False):
    if fused_modconv:
        raise NotImplementedError("fused_modconv=True is the reference's grouped branch (dead in training, "
                                  "stylegan.py:437-444); call with fused_modconv=False")
    n = x.shape[0]
    dcoefs = demod_coefficients(weight, styles) if demodulate else None
    x = x * styles.to(x.dtype).reshape(n, -1, 1, 1, 1)
    x = conv3d_resample(x, weight.to(x.dtype), f=resample_filter, up=up, down=down, padding=padding,
                        flip_weight=flip_weight)
    if dcoefs is not None:
        d = dcoefs.to(x.dtype).reshape(n, -1, 1, 1, 1)
        return torch.addcmul(noise.to(x.dtype), x, d) if noise is not None else x * d
    return x + noise.to(x.dtype) if noise is not None else x


def synthesis_conv(x, weight, styles, noise_field, noise_strength, bias, up=1, padding=1, resample_filter=None,
                   act="lrelu", demodulate=True):
    """One SynthesisLayer / OutBlock body (stylegan.py:526-531, 542-546): modulated conv, noise
    `noise_strength * noise_field` (field [N|1,1,R,R,R] or None), bias, activation.  The style
    scaling is ONE prologue pass (b2c_modulate), demodulation + noise + bias + activation ONE epilogue
    pass (b2c_mod_epilogue); the small reductions of their gradients ride in the same backward passes.
    Falls back to the elementwise formulation when the channel count has no 16-byte vectors."""
    cin, cout = x.shape[1], weight.shape[0]
    cd = BF.compute_dtype(x)
    if not (BF.mod_supported(cin, cd) and BF.mod_supported(cout, cd)):
        nz = None if noise_field is None else noise_field * noise_strength
        y = modulated_conv3d(x=x, weight=weight, styles=styles, noise=nz, up=up, padding=padding,
                             resample_filter=resample_filter, demodulate=demodulate, flip_weight=(up == 1),
                             fused_modconv=False)
        return bias_act(y, bias, act=act)
    dcoefs = demod_coefficients(weight, styles) if demodulate else None
    xs = BF.modulate(x, styles)
    y = conv3d_resample(xs, weight, f=resample_filter, up=up, padding=padding, flip_weight=(up == 1))
    a, slope = _ACT[act]
    return BF.mod_epilogue(y, dcoefs, noise_field, noise_strength, bias, act=a, slope=slope)


def bias_act(x, b=None, dim=1, act="linear"):
    """stylegan.py:118-124 (lrelu slope 0.2, no gain)."""
    if b is not None:
        x = x + b.to(x.dtype).reshape([-1 if i == dim else 1 for i in range(x.ndim)])
    if act == "linear":
        return x
    if act == "lrelu":
        return BF.activation(x, "lrelu", 0.2)
    if act == "relu":
        return BF.activation(x, "relu")
    if act == "tanh":
        return BF.activation(x, "tanh")
    raise NotImplementedError("activation %r is not used by the reference's networks" % (act,))
