"""ORACLE (test infrastructure, never on the product path): CPU restatement, from stock PyTorch
ops, of the StyleGAN2 convolution funnel of the reference for the ISOTROPIC arguments it uses
(one int for up / down / padding; 3-D resample filter):

  upfirdn3d          <- /root/reference/stylegan.py:172-200  (_upfirdn3d_ref)
  conv3d_wrapper     <- /root/reference/stylegan.py:164-170  (_conv3d_wrapper)
  conv3d_resample    <- /root/reference/stylegan.py:202-294
  modulated_conv3d   <- /root/reference/stylegan.py:396-444  (live branch: fused_modconv=False)
  bias_act           <- /root/reference/stylegan.py:118-124
  make_filter        <- /root/reference/stylegan.py:462-465  (the odd 4x4x4 [1,3,3,1] filter)

Quirks kept on purpose: FIR gain `up**2` (not up**3) after an up-sampling
conv, the non-separable 4x4x4 filter, lrelu slope 0.2 without gain.

Pinned: oracle/gen_golden_stylegan.py runs the UNMODIFIED reference functions next to these on
seeded inputs in the build container, asserts equality there and stores the vectors in
tests/golden/stylegan_ops_ref.npz; tests/test_oracle_models.py replays them.
"""
import torch
import torch.nn.functional as F


def make_filter(taps=(1, 3, 3, 1)):
    f = torch.Tensor(list(taps))
    f = f[None, None, :] * f[None, :, None]
    f = f.repeat((1, 4, 1)).reshape(4, 4, 4)
    return f / f.sum()


def upfirdn3d(x, f, up=1, down=1, pad=(0, 0), flip_filter=False, gain=1):
    """pad = (before, after), the same for the three axes."""
    if f is None:
        f = torch.ones([1, 1, 1], dtype=x.dtype, device=x.device)
    b, c, d, h, w = x.shape
    if up > 1:                                           # zero insertion AFTER each sample
        z = x.new_zeros(b, c, d, up, h, up, w, up)
        z[:, :, :, 0, :, 0, :, 0] = x
        x = z.reshape(b, c, d * up, h * up, w * up)
    p0, p1 = pad
    x = F.pad(x, [max(p0, 0), max(p1, 0)] * 3)
    c0, c1 = max(-p0, 0), max(-p1, 0)
    x = x[:, :, c0:x.shape[2] - c1, c0:x.shape[3] - c1, c0:x.shape[4] - c1]
    f = f * (gain ** (f.ndim / 3))
    if not flip_filter:
        f = f.flip(list(range(f.ndim)))
    k = f[None, None].repeat([c, 1, 1, 1, 1]).to(dtype=x.dtype, device=x.device)
    x = F.conv3d(x, k, groups=c)
    return x[:, :, ::down, ::down, ::down]


def conv3d_wrapper(x, w, stride=1, padding=0, transpose=False, flip_weight=True):
    if not flip_weight:
        w = w.flip([2, 3, 4])
    op = F.conv_transpose3d if transpose else F.conv3d
    return op(x, w, stride=stride, padding=padding)


def conv3d_resample(x, w, f=None, up=1, down=1, padding=0, flip_weight=True, flip_filter=False):
    k = int(w.shape[2])
    fk = 1 if f is None else int(f.shape[0])
    p0 = p1 = int(padding)
    if up > 1:
        p0 += (fk + up - 1) // 2
        p1 += (fk - up) // 2
    if down > 1:
        p0 += (fk - down + 1) // 2
        p1 += (fk - down) // 2
    if k == 1 and down > 1 and up == 1:
        x = upfirdn3d(x, f, down=down, pad=(p0, p1), flip_filter=flip_filter)
        return conv3d_wrapper(x, w, flip_weight=flip_weight)
    if k == 1 and up > 1 and down == 1:
        x = conv3d_wrapper(x, w, flip_weight=flip_weight)
        return upfirdn3d(x, f, up=up, pad=(p0, p1), gain=up ** 2, flip_filter=flip_filter)
    if down > 1 and up == 1:
        x = upfirdn3d(x, f, pad=(p0, p1), flip_filter=flip_filter)
        return conv3d_wrapper(x, w, stride=down, flip_weight=flip_weight)
    if up > 1:
        p0 -= k - 1
        p1 -= k - up
        pt = max(min(-p0, -p1), 0)
        x = conv3d_wrapper(x, w.transpose(0, 1), stride=up, padding=pt, transpose=True, flip_weight=not flip_weight)
        x = upfirdn3d(x, f, pad=(p0 + pt, p1 + pt), gain=up ** 2, flip_filter=flip_filter)
        if down > 1:
            x = upfirdn3d(x, f, down=down, flip_filter=flip_filter)
        return x
    if p0 == p1 and p0 >= 0:
        return conv3d_wrapper(x, w, padding=p0, flip_weight=flip_weight)
    x = upfirdn3d(x, None, pad=(p0, p1), gain=up ** 2, flip_filter=flip_filter)
    return conv3d_wrapper(x, w, flip_weight=flip_weight)


def modulated_conv3d(x, weight, styles, noise=None, up=1, down=1, padding=0, resample_filter=None,
                     demodulate=True, flip_weight=True, fused_modconv=False):
    if fused_modconv:
        raise NotImplementedError("the grouped fused branch is dead code in training (stylegan.py:437-444)")
    n = x.shape[0]
    dcoefs = None
    if demodulate:
        ws = weight.unsqueeze(0) * styles.reshape(n, 1, -1, 1, 1, 1)
        dcoefs = (ws.square().sum(dim=[2, 3, 4, 5]) + 1e-8).rsqrt()
    x = x * styles.to(x.dtype).reshape(n, -1, 1, 1, 1)
    x = conv3d_resample(x, weight.to(x.dtype), f=resample_filter, up=up, down=down, padding=padding,
                        flip_weight=flip_weight)
    if demodulate and noise is not None:
        return torch.addcmul(noise.to(x.dtype), x, dcoefs.to(x.dtype).reshape(n, -1, 1, 1, 1))
    if demodulate:
        return x * dcoefs.to(x.dtype).reshape(n, -1, 1, 1, 1)
    if noise is not None:
        return x + noise.to(x.dtype)
    return x


def bias_act(x, b=None, act="linear"):
    if b is not None:
        x = x + b.reshape(1, -1, 1, 1, 1)
    if act == "lrelu":
        return F.leaky_relu(x, 0.2)
    if act == "relu":
        return F.relu(x)
    if act == "tanh":
        return torch.tanh(x)
    return x


# the call signatures the reference's layers issue (stylegan.py:476-487, 521-532, 542-546, 758-790)
CASES = [
    # name, kind, cin, cout, k, up, down, res, demodulate, noise
    ("synth_conv1", "mod", 8, 8, 3, 1, 1, 8, True, True),          # SynthesisLayer conv1
    ("synth_conv0_up", "mod", 16, 8, 3, 2, 1, 4, True, True),      # SynthesisLayer conv0 (up=2)
    ("torgb", "mod", 8, 1, 1, 1, 1, 8, False, False),              # OutBlock 1x1x1, no demodulation
    ("disc_conv0", "plain", 8, 8, 3, 1, 1, 8, None, None),         # DiscriminatorBlock conv0
    ("disc_conv1_down", "plain", 8, 16, 3, 1, 2, 8, None, None),   # DiscriminatorBlock conv1 (down=2)
    ("disc_skip_down", "plain", 8, 16, 1, 1, 2, 8, None, None),    # DiscriminatorBlock skip (1x1x1, down=2)
    ("fromrgb", "plain", 1, 8, 1, 1, 1, 8, None, None),            # fromrgb 1x1x1
    ("skip_up_1x1", "plain", 16, 8, 1, 2, 1, 4, None, None),       # SynthesisBlock resnet skip (1x1x1, up=2)
]


def case_inputs(name, kind, cin, cout, k, up, down, res, demod, noise, batch=2):
    g = torch.Generator().manual_seed(sum(map(ord, name)))
    x = torch.randn(batch, cin, res, res, res, generator=g)
    w = torch.randn(cout, cin, k, k, k, generator=g) / (cin * k ** 3) ** 0.5
    s = torch.randn(batch, cin, generator=g) * 0.5 + 1.0
    ro = res * up // down
    nz = torch.randn(batch, 1, ro, ro, ro, generator=g) * 0.1 if noise else None
    gy = torch.randn(batch, cout, ro, ro, ro, generator=g)
    return x, w, s, nz, gy


def run_case(ops, case, x, w, s, nz):
    """ops: a module / namespace with conv3d_resample and modulated_conv3d of the reference's signature."""
    name, kind, cin, cout, k, up, down, res, demod, noise = case
    f = make_filter()
    if kind == "mod":
        return ops.modulated_conv3d(x=x, weight=w, styles=s, noise=nz, up=up, padding=k // 2, resample_filter=f,
                                    demodulate=demod, flip_weight=(up == 1), fused_modconv=False)
    return ops.conv3d_resample(x=x, w=w, f=f, up=up, down=down, padding=k // 2, flip_weight=(up == 1))
