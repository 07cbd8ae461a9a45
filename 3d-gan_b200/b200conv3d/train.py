"""One GAN training step exactly as the reference trainer schedules it
(trainer.py:208-252 D_step, trainer.py:254-283 G_step, trainer.py:296-303 loop):
parameter freeze / unfreeze decides which of fwd / dgrad / wgrad run, Adam
betas=(0, 0.9) (trainer.py:77-78), WGAN or hinge loss, mixed precision through
autocast.  The reference asks for fp16 autocast + GradScaler (trainer.py:14,80-81);
the B200 path computes in bf16, where loss scaling is the identity, so the scaler
is constructed disabled.

Works with any pair of nn.Modules (the drop-in layers on CUDA, or stock torch layers
on the CPU for the reference arm of bench.py).  `reducer_d` / `reducer_g` are
optional :class:`b200conv3d.parallel.GradReducer`s for data-parallel training.
"""
import contextlib

import torch
import torch.nn as tnn
import torch.nn.utils.parametrize  # noqa: F401  (tnn.utils.parametrize.is_parametrized)
import torch.optim as optim


def per_sample_stateless(netD):
    """True when one D pass over cat(real, fake) is the same arithmetic as the two passes of trainer.py:228-241: no
    layer couples the samples of a batch (BatchNorm, minibatch-std), draws random numbers or changes state per call
    (dropout, random crops, noise injection, spectral norm - one power iteration per forward).  An allow-list: every
    leaf module must be a layer known to be per-sample and stateless, without buffers or parametrisations."""
    from . import nn as BN
    ok = (tnn.Conv3d, tnn.ConvTranspose3d, tnn.Linear, tnn.LayerNorm, tnn.LeakyReLU, tnn.ReLU, tnn.Tanh, tnn.Sigmoid,
          tnn.Identity, tnn.Flatten, BN.FusedAway)
    if any(tnn.utils.parametrize.is_parametrized(m) for m in netD.modules()):
        return False
    leaves = [m for m in netD.modules() if not list(m.children())]
    return all(isinstance(m, ok) and not hasattr(m, "weight_u") and not hasattr(m, "weight_orig") and
               not list(m.buffers(recurse=False)) for m in leaves)


class GanStep:
    def __init__(self, netG, netD, z_size=512, lrG=5e-5, lrD=1e-4, hinge=False, iterD=2,
                 autocast_dtype=torch.bfloat16, reducer_g=None, reducer_d=None, seed=None, batch_real_fake=True):
        from . import functional as BF
        BF.watch_optimizers()          # fused Adam does not bump Tensor._version (packed-filter cache)
        self.netG, self.netD = netG, netD
        self.z_size, self.hinge, self.iterD = z_size, hinge, iterD
        self.device = next(netG.parameters()).device
        fused = self.device.type == "cuda"
        self.optimizerD = optim.Adam(netD.parameters(), lr=lrD, betas=(0.0, 0.9), fused=fused)
        self.optimizerG = optim.Adam(netG.parameters(), lr=lrG, betas=(0.0, 0.9), fused=fused)
        self.autocast_dtype = autocast_dtype
        self.reducer_g, self.reducer_d = reducer_g, reducer_d
        self.batch_real_fake = bool(batch_real_fake) and self.device.type == "cuda" and per_sample_stateless(netD)
        self.gen = None
        if seed is not None:
            self.gen = torch.Generator(device=self.device)
            self.gen.manual_seed(seed)
        for p in list(netD.parameters()) + list(netG.parameters()):   # trainer.py:290-293
            p.requires_grad = False

    def _amp(self):
        if self.autocast_dtype is None or self.device.type != "cuda":
            return contextlib.nullcontext()
        return torch.autocast("cuda", dtype=self.autocast_dtype)

    def noise(self, n):
        return torch.randn(n, self.z_size, 1, 1, 1, dtype=torch.float, device=self.device,
                           generator=self.gen)

    def D_step(self, real):
        """trainer.py:208-252.  Returns (errD_real, errD_fake) as 0-d device tensors."""
        for p in self.netD.parameters():
            p.requires_grad = True
        self.netD.zero_grad()
        if self.reducer_d is not None:
            self.reducer_d.begin()
        with self._amp():
            fake = self.netG(self.noise(real.shape[0]))
            if self.batch_real_fake:
                # netD has no batch-coupled layer (LayerNorm / spectral norm only): D(real) and D(fake) of
                # trainer.py:222-241 as ONE pass over the concatenated batch - the same per-sample arithmetic, half
                # the launches, one weight-gradient GEMM per layer with twice the reduction and no accumulate pass
                n = real.shape[0]
                out = self.netD(torch.cat([real.to(fake.dtype), fake]))
                o_real, o_fake = out[:n], out[n:]
            else:
                o_real, o_fake = self.netD(real), self.netD(fake)
            if self.hinge:
                errD_real = tnn.functional.relu(1.0 - o_real).mean()
                errD_fake = tnn.functional.relu(1.0 + o_fake).mean()
                errD = errD_fake + errD_real
            else:
                errD_real = o_real.mean()
                errD_fake = o_fake.mean()
                errD = errD_fake - errD_real
            errD.backward()
        if self.reducer_d is not None:
            self.reducer_d.step(self.optimizerD)      # the update follows the averaged buckets one by one
        else:
            self.optimizerD.step()
        for p in self.netD.parameters():
            p.requires_grad = False
        return errD_real.detach(), errD_fake.detach()

    def G_step(self, batch):
        """trainer.py:254-283.  Returns (errG, fake)."""
        for p in self.netG.parameters():
            p.requires_grad = True
        self.netG.zero_grad()
        if self.reducer_g is not None:
            self.reducer_g.begin()
        with self._amp():
            fake = self.netG(self.noise(batch))
            errG = -self.netD(fake).mean()
            errG.backward()
        if self.reducer_g is not None:
            self.reducer_g.step(self.optimizerG)
        else:
            self.optimizerG.step()
        for p in self.netG.parameters():
            p.requires_grad = False
        return errG.detach(), fake.detach()

    def step(self, reals):
        """One iteration of trainer.py:296-303: iterD discriminator steps, one generator step.
        `reals` is an iterable yielding iterD real batches [B,1,128,128,128] on the device."""
        out = []
        batch = None
        for real in reals:
            batch = real.shape[0]
            out.append(self.D_step(real))
        errG, fake = self.G_step(batch)
        return out, errG, fake


class DevicePrefetcher:
    """Host -> device input pipeline for the step (trainer.py:84,300 `data.to(device)`): yields
    device tensors while the NEXT pinned host batch is already being copied on a side stream,
    so the copy overlaps the running D_step instead of stalling it."""

    def __init__(self, batches, device):
        self.it = iter(batches)
        self.device = torch.device(device)
        self.stream = torch.cuda.Stream(self.device)
        self._next = None
        self._preload()

    def _preload(self):
        try:
            host = next(self.it)
        except StopIteration:
            self._next = None
            return
        with torch.cuda.stream(self.stream):
            self._next = host.to(self.device, non_blocking=True)
            self._ready = torch.cuda.Event()
            self._ready.record(self.stream)

    def __iter__(self):
        return self

    def __next__(self):
        if self._next is None:
            raise StopIteration
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(self._ready)
        out = self._next
        out.record_stream(cur)
        self._preload()
        return out
