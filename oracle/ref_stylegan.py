"""ORACLE (test infrastructure, never on the product path): CPU restatement, from stock PyTorch
ops, of the reference's StyleGAN networks and losses.

  dense / Mapping             <- /root/reference/stylegan.py:298-392
  ConvLayer                   <- /root/reference/stylegan.py:446-487
  SynthLayer / ToImage / SynthBlock / Synthesis / Generator
                              <- /root/reference/stylegan.py:489-718
  DiscBlock / minibatch_std / DiscEpilogue / Discriminator
                              <- /root/reference/stylegan.py:721-928
  adain / StyleGen1           <- /root/reference/stylegan.py:931-1149
  step_D / step_G             <- /root/reference/stylegan.py:37-100 (StyleGAN2Loss)

Attribute names equal the reference's so that state_dicts interchange (`synthesis.b8.conv0.affine.weight`,
`b16.fromrgb.weight`, `latentMapping.0.weight`, ...).  Only the live configuration is restated:
'skip' synthesis, 'resnet' discriminator, `fused_modconv=False`, no fp16 blocks, no labels.

Pinned: oracle/gen_golden_stylegan_net.py runs the UNMODIFIED reference classes next to these on seeded
inputs in the build container, asserts equality there and stores the vectors in
tests/golden/stylegan_net_ref.npz; tests/test_stylegan_net.py replays them.
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

import ref_stylegan_ops as O


class Dense(nn.Module):
    def __init__(self, fin, fout, act="linear", lr_mul=1.0, bias_init=0.0):
        super().__init__()
        self.act, self.wg, self.bg = act, lr_mul / math.sqrt(fin), lr_mul
        self.weight = nn.Parameter(torch.randn(fout, fin) / lr_mul)
        self.bias = nn.Parameter(torch.full([fout], float(bias_init)))

    def forward(self, x):
        y = F.linear(x, self.weight * self.wg, self.bias * self.bg)
        return F.leaky_relu(y, 0.2) if self.act == "lrelu" else y


class Mapping(nn.Module):
    def __init__(self, z_dim, w_dim, num_ws, layers=8):
        super().__init__()
        self.n, self.num_ws = layers, num_ws
        for i in range(layers):
            setattr(self, "fc%d" % i, Dense(z_dim if i == 0 else w_dim, w_dim, "lrelu", 0.01))
        self.register_buffer("w_avg", torch.zeros(w_dim))

    def forward(self, z, **_):
        x = z.squeeze().float()
        x = x * (x.square().mean(1, keepdim=True) + 1e-8).rsqrt()
        for i in range(self.n):
            x = getattr(self, "fc%d" % i)(x)
        return x.unsqueeze(1).repeat(1, self.num_ws, 1)


class ConvLayer(nn.Module):
    def __init__(self, cin, cout, k, bias=True, act="linear", up=1, down=1):
        super().__init__()
        self.act, self.up, self.down, self.k = act, up, down, k
        self.gain = 1.0 / math.sqrt(cin * k ** 3)
        self.weight = nn.Parameter(torch.randn(cout, cin, k, k, k))
        self.bias = nn.Parameter(torch.zeros(cout)) if bias else None

    def forward(self, x):
        y = O.conv3d_resample(x, self.weight * self.gain, f=O.make_filter(), up=self.up, down=self.down,
                              padding=self.k // 2, flip_weight=(self.up == 1))
        return O.bias_act(y, self.bias, self.act)


class SynthLayer(nn.Module):
    def __init__(self, cin, cout, w_dim, res, up=1):
        super().__init__()
        self.res, self.up = res, up
        self.affine = Dense(w_dim, cin, bias_init=1.0)
        self.weight = nn.Parameter(torch.randn(cout, cin, 3, 3, 3))
        self.register_buffer("noise_const", torch.randn(res, res))
        self.noise_strength = nn.Parameter(torch.zeros([]))
        self.bias = nn.Parameter(torch.zeros(cout))

    def forward(self, x, w, noise_mode="random", noise=None):
        if noise_mode == "random":
            r = self.res
            nz = (torch.randn(x.shape[0], 1, r, r, r, device=x.device) if noise is None else noise) * self.noise_strength
        elif noise_mode == "const":
            nz = self.noise_const * self.noise_strength
        else:
            nz = None
        y = O.modulated_conv3d(x, self.weight, self.affine(w), noise=nz, up=self.up, padding=1,
                               resample_filter=O.make_filter(), flip_weight=(self.up == 1))
        return O.bias_act(y, self.bias, "lrelu")


class ToImage(nn.Module):
    def __init__(self, cin, cimg, w_dim):
        super().__init__()
        self.gain = 1.0 / math.sqrt(cin)
        self.affine = Dense(w_dim, cin, bias_init=1.0)
        self.weight = nn.Parameter(torch.randn(cimg, cin, 1, 1, 1))
        self.bias = nn.Parameter(torch.zeros(cimg))

    def forward(self, x, w):
        y = O.modulated_conv3d(x, self.weight, self.affine(w) * self.gain, demodulate=False)
        return O.bias_act(y, self.bias)


class SynthBlock(nn.Module):
    def __init__(self, cin, cout, w_dim, res):
        super().__init__()
        self.cin = cin
        if cin == 0:
            self.const = nn.Parameter(torch.randn(cout, res, res, res))
        else:
            self.conv0 = SynthLayer(cin, cout, w_dim, res, up=2)
        self.conv1 = SynthLayer(cout, cout, w_dim, res)
        self.torgb = ToImage(cout, 1, w_dim)
        self.num_conv = 1 if cin == 0 else 2

    def forward(self, x, ws, img, noises=None, **kw):
        w = list(ws.unbind(1))
        nz = list(noises) if noises is not None else [None, None]
        if self.cin == 0:
            x = self.conv1(self.const.unsqueeze(0).repeat(ws.shape[0], 1, 1, 1, 1), w[0], noise=nz[0], **kw)
        else:
            x = self.conv0(x, w[0], noise=nz[0], **kw)
            x = self.conv1(x, w[1], noise=nz[1], **kw)
        if img is not None:
            img = O.upfirdn3d(img, O.make_filter(), up=2, pad=(2, 1), gain=8)
        y = self.torgb(x, w[self.num_conv])
        return x, (y if img is None else img + y)


class Synthesis(nn.Module):
    def __init__(self, w_dim, img_resolution=128, channel_base=4096, channel_max=512):
        super().__init__()
        self.res = [2 ** i for i in range(2, int(math.log2(img_resolution)) + 1)]
        ch = {r: min(channel_base // r, channel_max) for r in self.res}
        self.num_ws = 0
        for r in self.res:
            b = SynthBlock(ch[r // 2] if r > 4 else 0, ch[r], w_dim, r)
            self.num_ws += b.num_conv + (1 if r == img_resolution else 0)
            setattr(self, "b%d" % r, b)

    def forward(self, ws, noises=None, **kw):
        x = img = None
        at = 0
        for i, r in enumerate(self.res):
            b = getattr(self, "b%d" % r)
            x, img = b(x, ws[:, at:at + b.num_conv + 1].float(), img, noises=None if noises is None else noises[i], **kw)
            at += b.num_conv
        return torch.tanh(img)


class Generator(nn.Module):
    def __init__(self, z_dim, channel_base, w_dim=512, img_resolution=128):
        super().__init__()
        self.synthesis = Synthesis(w_dim, img_resolution, channel_base)
        self.mapping = Mapping(z_dim, w_dim, self.synthesis.num_ws)

    def forward(self, z, **kw):
        ws = self.mapping(z)
        return self.synthesis(ws, **kw), ws


class DiscBlock(nn.Module):
    def __init__(self, cin, ctmp, cout):
        super().__init__()
        self.cin = cin
        if cin == 0:
            self.fromrgb = ConvLayer(1, ctmp, 1, act="lrelu")
        self.conv0 = ConvLayer(ctmp, ctmp, 3, act="lrelu")
        self.conv1 = ConvLayer(ctmp, cout, 3, act="lrelu", down=2)
        self.skip = ConvLayer(ctmp, cout, 1, bias=False, down=2)

    def forward(self, x, img):
        if self.cin == 0:
            x = self.fromrgb(img)
        return self.skip(x) + self.conv1(self.conv0(x))


def minibatch_std(x, group=4):
    n, c = x.shape[:2]
    g = min(group, n)
    y = x.reshape(g, -1, 1, c, *x.shape[2:])
    y = ((y - y.mean(0)).square().mean(0) + 1e-8).sqrt().mean(dim=[2, 3, 4, 5])
    y = y.reshape(-1, 1, 1, 1, 1).repeat(g, 1, *x.shape[2:])
    return torch.cat([x, y], 1)


class DiscEpilogue(nn.Module):
    def __init__(self, c, res=4):
        super().__init__()
        self.conv = ConvLayer(c + 1, c, 3, act="lrelu")
        self.fc = Dense(c * res ** 3, c, "lrelu")
        self.out = Dense(c, 1)

    def forward(self, x):
        return self.out(self.fc(self.conv(minibatch_std(x)).flatten(1)))


class Discriminator(nn.Module):
    def __init__(self, channel_base, img_resolution=128, channel_max=512):
        super().__init__()
        self.res = [2 ** i for i in range(int(math.log2(img_resolution)), 2, -1)]
        ch = {r: min(channel_base // r, channel_max) for r in self.res + [4]}
        for r in self.res:
            setattr(self, "b%d" % r, DiscBlock(ch[r] if r < img_resolution else 0, ch[r], ch[r // 2]))
        self.b4 = DiscEpilogue(ch[4])

    def forward(self, img):
        x = None
        for r in self.res:
            x = getattr(self, "b%d" % r)(x, img)
        return self.b4(x)


# ------------------------------------------------------------------ v1-like generator
def adain(f, style):
    c = f.shape[1]
    flat = f.reshape(f.shape[0], c, -1)
    std = (flat.var(dim=2) + 1e-5).sqrt()[:, :, None, None, None]
    mean = flat.mean(dim=2)[:, :, None, None, None]
    return (f - mean) / std * style[:, c:, None, None, None] + style[:, :c, None, None, None]


class StyleGen1(nn.Module):
    W = [512, 512, 256, 256, 128, 128, 64, 64, 32, 32, 16]

    def __init__(self, nz=512, mixing=True):
        super().__init__()
        self.mixing = mixing
        seq = []
        for _ in range(8):
            seq += [nn.Linear(nz, nz), nn.LeakyReLU(0.2)]
        self.latentMapping = nn.Sequential(*seq)
        for i in range(10):
            setattr(self, "C%d" % (i + 1), nn.Sequential(nn.Conv3d(self.W[i], self.W[i + 1], 3, 1, 1, bias=False),
                                                         nn.LeakyReLU(0.2)))
        self.C_out = nn.Sequential(nn.Conv3d(16, 1, 3, 1, 1, bias=False), nn.Tanh())
        for i in range(11):
            setattr(self, "A%d" % (i + 1), nn.Linear(512, 2 * self.W[i]))

    def forward(self, z):
        mix = self.mixing and self.training
        swap = torch.randint(6, (1, 1, 1)).item() if mix else -1
        w = self.latentMapping(z.squeeze())
        h = adain(torch.ones(z.shape[0], 512, 4, 4, 4, dtype=z.dtype, device=z.device), self.A1(w))
        point = 0
        for i in range(1, 11):
            h = getattr(self, "C%d" % i)(h)
            if mix and (i % 2 == 1 or i == 10):
                if point == swap:
                    w = w[torch.randperm(w.shape[0])]
                point += 1
            h = adain(h, getattr(self, "A%d" % (i + 1))(w))
            if i % 2 == 1:
                h = F.interpolate(h, scale_factor=2, mode="trilinear")
        return self.C_out(h)


# ------------------------------------------------------------------ losses
def step_D(G, D, step, real, z, r1_gamma=10.0, v1=False, **gkw):
    """Accumulates the discriminator gradients of one StyleGAN2Loss.step_D; returns (loss_real, loss_gen)."""
    fake = G(z) if v1 else G(z, **gkw)[0]
    l_gen = F.softplus(D(fake)).mean()
    l_gen.backward()
    r1 = (step % 16 == 0) or v1
    x = real.detach().requires_grad_(r1)
    logits = D(x)
    pen = 0
    if r1:
        (g,) = torch.autograd.grad(logits.sum(), x, retain_graph=True)
        pen = g.square().sum([1, 2, 3]) * (r1_gamma / 2)
    l_real = (logits * 0 + F.softplus(-logits) + pen).mean()
    l_real.backward()
    return l_real.detach(), l_gen.detach()


def step_G(G, D, step, z, pl_mean, pl_noise=None, pl_weight=2.0, pl_decay=0.01, **gkw):
    """One StyleGAN2Loss.step_G for the StyleGAN2 generator; `pl_mean` is updated in place."""
    img, _ = G(z, **gkw)
    loss = F.softplus(-D(img)).mean()
    loss.backward()
    if step % 16 == 0:
        img, ws = G(z[:z.shape[0] // 2], **gkw)
        nz = (torch.randn_like(img) if pl_noise is None else pl_noise) / math.sqrt(img.shape[2] * img.shape[3])
        (g,) = torch.autograd.grad((img * nz).sum(), ws, retain_graph=True)
        lengths = g.square().sum(2).mean(1).sqrt()
        mean = pl_mean.lerp(lengths.mean(), pl_decay)
        pl_mean.copy_(mean.detach())
        pl = (img[:, 0, 0, 0, 0] * 0 + (lengths - mean).square() * pl_weight).mean()
        pl.backward()
        loss = pl + loss
    return loss.detach(), img
