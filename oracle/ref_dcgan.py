"""ORACLE (test infrastructure, never on the product path): a CPU restatement of the
reference's DCGAN training step from stock PyTorch layers.

The reference holds no arithmetic of its own on this path: its
Generator / Discriminator are nn.Sequential tables of torch layers and its trainer
calls them.  This file restates those tables and the two trainer phases so that they
can run on the GPU box, where /root/reference does not exist:

  Generator      <- /root/reference/dcgan.py:45-70, init 74-81
  Discriminator  <- /root/reference/dcgan.py:158-182 (default LayerNorm branch), init 184-191
  d_step / g_step<- /root/reference/trainer.py:208-252 / 254-283 (WGAN branch, no GP)
  Adam betas     <- /root/reference/trainer.py:77-78

Pinned: tests/test_oracle_models.py loads the golden vectors in tests/golden/dcgan_*.npz,
which oracle/gen_golden.py produced by importing the UNMODIFIED reference modules from
/root/reference in the build container (torch version recorded in the file), and checks
this restatement against them bit-for-bit on the same torch build.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
import this file.
"""
import torch
import torch.nn as nn
import torch.optim as optim


def _init(m):
    name = m.__class__.__name__
    if name.find("Conv") != -1:
        nn.init.normal_(m.weight.data, 0.0, 0.02)
    elif name.find("BatchNorm") != -1:
        nn.init.normal_(m.weight.data, 1.0, 0.02)
        nn.init.constant_(m.bias.data, 0)


class Generator(nn.Module):
    def __init__(self, z_size=512, filterG=128):
        super().__init__()
        nz, ngf = z_size, filterG
        self.main = nn.Sequential(
            nn.ConvTranspose3d(nz, ngf * 16, 4, stride=1), nn.BatchNorm3d(ngf * 16), nn.ReLU(True),
            nn.ConvTranspose3d(ngf * 16, ngf * 8, 4, stride=2, padding=1), nn.BatchNorm3d(ngf * 8), nn.ReLU(True),
            nn.ConvTranspose3d(ngf * 8, ngf * 4, 4, stride=2, padding=1), nn.BatchNorm3d(ngf * 4), nn.ReLU(True),
            nn.ConvTranspose3d(ngf * 4, ngf * 2, 4, stride=2, padding=1), nn.BatchNorm3d(ngf * 2), nn.ReLU(True),
            nn.ConvTranspose3d(ngf * 2, ngf, 4, stride=2, padding=1), nn.BatchNorm3d(ngf), nn.ReLU(True),
            nn.ConvTranspose3d(ngf, 1, 4, stride=2, padding=1), nn.Tanh())
        self.apply(_init)

    def forward(self, z):
        return self.main(z)


class Discriminator(nn.Module):
    def __init__(self, filterD=128):
        super().__init__()
        ndf = filterD
        self.main = nn.Sequential(
            nn.Conv3d(1, ndf, 4, stride=2, padding=1, bias=False), nn.LayerNorm([ndf, 64, 64, 64]),
            nn.LeakyReLU(0.2, inplace=True),
            nn.Conv3d(ndf, ndf * 2, 4, stride=2, padding=1, bias=False), nn.LayerNorm([ndf * 2, 32, 32, 32]),
            nn.LeakyReLU(0.2, inplace=True),
            nn.Conv3d(ndf * 2, ndf * 4, 4, stride=2, padding=1, bias=False), nn.LayerNorm([ndf * 4, 16, 16, 16]),
            nn.LeakyReLU(0.2, inplace=True),
            nn.Conv3d(ndf * 4, ndf * 8, 4, stride=2, padding=1, bias=False), nn.LayerNorm([ndf * 8, 8, 8, 8]),
            nn.LeakyReLU(0.2, inplace=True),
            nn.Conv3d(ndf * 8, ndf * 16, 4, stride=2, padding=1, bias=False), nn.LayerNorm([ndf * 16, 4, 4, 4]),
            nn.LeakyReLU(0.2, inplace=True),
            nn.Conv3d(ndf * 16, 1, (4, 4, 4), stride=1, padding=0, bias=False))
        self.apply(_init)

    def forward(self, x):
        return self.main(x)


class RefTrainer:
    """The two phases of trainer.Trainer for --dcgan (WGAN loss, trainer.py:235-247) or any
    model with --hinge (trainer.py:222-233), fp32, any device."""

    def __init__(self, netG, netD, z_size=512, lrG=5e-5, lrD=1e-4, hinge=False):
        self.netG, self.netD, self.z_size, self.hinge = netG, netD, z_size, hinge
        self.device = next(netG.parameters()).device
        self.optimizerD = optim.Adam(netD.parameters(), lr=lrD, betas=(0.0, 0.9))
        self.optimizerG = optim.Adam(netG.parameters(), lr=lrG, betas=(0.0, 0.9))
        for p in list(netD.parameters()) + list(netG.parameters()):
            p.requires_grad = False

    def d_step(self, real, noise=None):
        for p in self.netD.parameters():
            p.requires_grad = True
        self.netD.zero_grad()
        if noise is None:
            noise = torch.randn(real.shape[0], self.z_size, 1, 1, 1, dtype=torch.float, device=self.device)
        fake = self.netG(noise)
        if self.hinge:
            errD_real = torch.relu(1.0 - self.netD(real)).mean()
            errD_fake = torch.relu(1.0 + self.netD(fake)).mean()
            errD = errD_fake + errD_real
        else:
            errD_real = self.netD(real).mean()
            errD_fake = self.netD(fake).mean()
            errD = errD_fake - errD_real
        errD.backward()
        self.optimizerD.step()
        for p in self.netD.parameters():
            p.requires_grad = False
        return errD_real.item(), errD_fake.item()

    def g_step(self, batch, noise=None):
        for p in self.netG.parameters():
            p.requires_grad = True
        self.netG.zero_grad()
        if noise is None:
            noise = torch.randn(batch, self.z_size, 1, 1, 1, dtype=torch.float, device=self.device)
        fake = self.netG(noise)
        errG = -self.netD(fake).mean()
        errG.backward()
        self.optimizerG.step()
        for p in self.netG.parameters():
            p.requires_grad = False
        return errG.item(), fake.detach()
