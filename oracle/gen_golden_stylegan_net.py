"""Generates tests/golden/stylegan_net_ref.npz by running the UNMODIFIED reference classes
(/root/reference/stylegan.py: Generator, Discriminator, styleGAN_gen, StyleGAN2Loss) on the CPU of
the build container, and checks oracle/ref_stylegan.py against them on the spot.

StyleGAN2 part: img_resolution 32, channel_base 128, w_dim 32, z 16, batch 4 (constructor arguments
of the reference's own classes - the Trainer hard-codes 128^3 / w_dim 512, too large for a fixture).
`noise_strength` parameters are set to 0.1 (they are initialised to 0, which would hide the noise
path).  Losses: the reference's StyleGAN2Loss with `style_mixing_prob=0` and the synthesis network
called with `noise_mode='const'` (both are arguments of the reference), steps 0 (R1 and path-length
terms active) under the Trainer's freeze pattern (trainer.py:208-283).

v1-like generator (styleGAN_gen): widths are hard-coded (512..16 channels, 128^3 output), batch 2;
eval-mode forward and a training-mode forward (batch-shuffle style mixing from the CPU RNG).

  PYTHONDONTWRITEBYTECODE=1 python oracle/gen_golden_stylegan_net.py
"""
import argparse
import functools
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ref_stylegan as R  # noqa: E402
from gen_golden import REF, stub_modules  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "stylegan_net_ref.npz")
CFG = dict(res=32, cb=128, w_dim=32, z=16, batch=4, seed=20261104)


def set_noise_strength(net, v=0.1):
    with torch.no_grad():
        for k, p in net.named_parameters():
            if k.endswith("noise_strength"):
                p.fill_(v)


def random_noises(net_res, batch, first=4):
    """The N(0,1) fields a noise_mode='random' forward draws, in execution order (stylegan.py:522)."""
    out = []
    for r in net_res:
        out.append([torch.randn(batch, 1, r, r, r) for _ in range(1 if r == first else 2)])
    return out


def close(a, b, what, rtol=2e-4, atol=1e-5):
    torch.testing.assert_close(a, b, rtol=rtol, atol=atol, msg=lambda m: "%s: %s" % (what, m))


def main():
    stub_modules()
    sys.path.insert(0, REF)
    sys.dont_write_bytecode = True
    import stylegan as S          # noqa: E402  (the reference, unmodified)
    c = CFG
    p = argparse.Namespace(z_size=c["z"], filterG=c["cb"], filterD=c["cb"], device="cpu", stylegan2=True, stylegan=False)
    out = {"torch_version": np.array(torch.__version__), **{"cfg." + k: np.array(v) for k, v in c.items()}}

    torch.manual_seed(c["seed"])
    D = S.Discriminator(p, img_resolution=c["res"])                      # trainer.py:59-60: D first
    G = S.Generator(p, w_dim=c["w_dim"], img_resolution=c["res"])
    torch.manual_seed(c["seed"])
    oD = R.Discriminator(c["cb"], img_resolution=c["res"])
    oG = R.Generator(c["z"], c["cb"], w_dim=c["w_dim"], img_resolution=c["res"])
    for ref, ora, pre in ((G, oG, "g0."), (D, oD, "d0.")):
        sd, so = ref.state_dict(), ora.state_dict()
        assert list(sd) == list(so), (set(sd) ^ set(so))
        for k in sd:
            assert torch.equal(sd[k], so[k]), k
            out[pre + k] = sd[k].numpy().copy()
    set_noise_strength(G), set_noise_strength(oG)

    torch.manual_seed(5)
    z = torch.randn(c["batch"], c["z"], 1, 1, 1)
    real = torch.rand(c["batch"], 1, c["res"], c["res"], c["res"]) * 2 - 1
    out["z"], out["real_sample"] = z.numpy(), real[:, 0, ::4, ::4, ::4].numpy()
    out["real_seeded_by"] = np.array("torch.manual_seed(5); z = randn(B,z,1,1,1); real = rand(B,1,R,R,R)*2-1")
    with torch.no_grad():
        img_c, ws = G(z, noise_mode="const")
        logits = D(img_c)
        torch.manual_seed(6)
        img_r, _ = G(z)                                                   # noise_mode='random' from the CPU RNG
        o_img_c, o_ws = oG(z, noise_mode="const")
        torch.manual_seed(6)
        o_img_r, _ = oG(z, noises=random_noises(oG.synthesis.res, c["batch"]))
        close(o_img_c, img_c, "img const"), close(o_ws, ws, "ws"), close(o_img_r, img_r, "img random")
        close(oD(img_c), logits, "logits")
    out.update(img_const=img_c[:, 0, ::2, ::2, ::2].numpy(), img_const_sum=np.array(img_c.double().sum().item()),
               ws=ws[:, 0].numpy(), img_random=img_r[:, 0, ::2, ::2, ::2].numpy(), logits=logits.numpy())

    # one D step and one G step of the reference loss at step 0, Trainer freeze pattern
    loss = S.StyleGAN2Loss(p, G.mapping, functools.partial(G.synthesis, noise_mode="const"), D, style_mixing_prob=0)
    for net in (G, D, oG, oD):
        for q in net.parameters():
            q.requires_grad = False
    opl = torch.zeros([])
    for q, oq in zip(D.parameters(), oD.parameters()):
        q.requires_grad = oq.requires_grad = True
    lr, lg = loss.step_D(0, real, z)
    olr, olg = R.step_D(oG, oD, 0, real, z, noise_mode="const")
    close(olr, lr, "D loss real"), close(olg, lg, "D loss gen")
    out["D_losses"] = np.array([lr.item(), lg.item()])
    for (k, q), oq in zip(D.named_parameters(), oD.parameters()):
        close(oq.grad, q.grad, "dgrad " + k, rtol=2e-3, atol=1e-6 + 2e-4 * q.grad.abs().max().item())
        out["dgrad." + k] = q.grad.numpy().copy()
        q.requires_grad = oq.requires_grad = False
        q.grad = oq.grad = None
    for q, oq in zip(G.parameters(), oG.parameters()):
        q.requires_grad = oq.requires_grad = True
    z2 = torch.randn(c["batch"], c["z"], 1, 1, 1, generator=torch.Generator().manual_seed(7))
    torch.manual_seed(8)          # the first draw of step_G is the path-length noise (stylegan.py:86)
    lG, _ = loss.step_G(0, z2)
    pl_noise = torch.randn(c["batch"] // 2, 1, c["res"], c["res"], c["res"], generator=torch.Generator().manual_seed(8))
    olG, _ = R.step_G(oG, oD, 0, z2, opl, pl_noise=pl_noise, noise_mode="const")
    close(olG, lG.detach(), "G loss"), close(opl, loss.pl_mean, "pl_mean")
    out.update(z2=z2.numpy(), G_loss=np.array(lG.item()), pl_mean=np.array(loss.pl_mean.item()),
               pl_noise_seeded_by=np.array("torch.manual_seed(8); randn(B/2,1,R,R,R) (global CPU generator)"))
    for (k, q), oq in zip(G.named_parameters(), oG.parameters()):
        if q.grad is None:
            assert oq.grad is None, k
            continue
        close(oq.grad, q.grad, "ggrad " + k, rtol=2e-3, atol=1e-6 + 2e-4 * q.grad.abs().max().item())
        out["ggrad." + k] = q.grad.numpy().copy()

    # ---- v1-like generator (full hard-coded widths)
    p1 = argparse.Namespace(z_size=512)
    torch.manual_seed(c["seed"] + 1)
    G1 = S.styleGAN_gen(p1)
    torch.manual_seed(c["seed"] + 1)
    oG1 = R.StyleGen1()
    assert all(torch.equal(a, b) for a, b in zip(G1.state_dict().values(), oG1.state_dict().values()))
    assert list(G1.state_dict()) == list(oG1.state_dict())
    z1 = torch.randn(2, 512, 1, 1, 1, generator=torch.Generator().manual_seed(9))
    G1.eval(), oG1.eval()
    with torch.no_grad():
        y_eval = G1(z1)
        close(oG1(z1), y_eval, "v1 eval")
    G1.train(), oG1.train()
    torch.manual_seed(10)
    y_tr = G1(z1)
    y_tr.square().mean().backward()
    torch.manual_seed(10)
    oy = oG1(z1)
    oy.square().mean().backward()
    close(oy, y_tr, "v1 train")
    out.update(v1_seed=np.array(c["seed"] + 1), v1_z=z1.numpy(), v1_eval=y_eval[:, 0, ::8, ::8, ::8].numpy(),
               v1_eval_sum=np.array(y_eval.double().sum().item()), v1_train=y_tr.detach()[:, 0, ::8, ::8, ::8].numpy())
    sd, so = dict(G1.named_parameters()), dict(oG1.named_parameters())
    for k in ("C1.0.weight", "C6.0.weight", "C10.0.weight", "C_out.0.weight", "A1.weight", "A11.weight", "latentMapping.0.weight"):
        g = sd[k].grad
        close(so[k].grad, g, "v1 grad " + k, rtol=2e-3, atol=1e-9 + 2e-4 * g.abs().max().item())
        flat = g.reshape(-1)
        out["v1grad." + k] = flat[:: max(1, flat.numel() // 4096)].numpy().copy()
        out["v1gradnorm." + k] = np.array(g.double().norm().item())
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes; D_losses", out["D_losses"], "G_loss", out["G_loss"],
          "pl_mean", out["pl_mean"], "keys", len(out))


if __name__ == "__main__":
    main()
