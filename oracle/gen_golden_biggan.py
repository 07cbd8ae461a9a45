"""Generates tests/golden/biggan_ref.npz by running the UNMODIFIED reference
(/root/reference: biggan.py, utils.py, trainer.py) on the CPU of the build container,
BigGAN-deep branch (`--biggan=True --hinge=True`, run.sh:1) at a reduced width.

Stubs as in oracle/gen_golden.py (modules without arithmetic).

  PYTHONDONTWRITEBYTECODE=1 python oracle/gen_golden_biggan.py
"""
import os
import sys
import tempfile

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from gen_golden import REF, params, stub_modules  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "biggan_ref.npz")


def main():
    stub_modules()
    sys.path.insert(0, REF)
    sys.dont_write_bytecode = True
    import trainer as ref_trainer          # noqa: E402  (the reference, unmodified)

    width, z, batch = 2, 8, 2
    p = params(None, width, z, batch)
    p.dcgan, p.biggan, p.hinge = False, True, True
    torch.manual_seed(20261104)
    with tempfile.TemporaryDirectory() as tmp:
        p.log_dir = tmp
        tr = ref_trainer.Trainer([torch.zeros(1)] * 8, p)
        g0 = {k: v.clone() for k, v in tr.netG.state_dict().items()}
        d0 = {k: v.clone() for k, v in tr.netD.state_dict().items()}
        torch.manual_seed(5)
        zin = torch.randn(batch, z, 1, 1, 1)
        real = torch.rand(batch, 1, 128, 128, 128) * 2 - 1
        # eval-mode probes do not touch the spectral-norm u / v buffers or BN running stats
        tr.netG.eval(), tr.netD.eval()
        with torch.no_grad():
            fake0 = tr.netG(zin)
            dreal0 = tr.netD(real)
        tr.netG.train(), tr.netD.train()
        for q in list(tr.netD.parameters()) + list(tr.netG.parameters()):
            q.requires_grad = False
        torch.manual_seed(77)
        tr.D_step(0, real)
        torch.manual_seed(78)
        tr.G_step(0)
        g1, d1 = tr.netG.state_dict(), tr.netD.state_dict()
        out = {
            "torch_version": np.array(torch.__version__), "width": width, "z": z, "batch": batch,
            "zin": zin.numpy(), "real_sample": real[:, 0, ::16, ::16, ::16].numpy(),
            "fake0_sample": fake0[:, 0, ::8, ::8, ::8].numpy(), "fake0_sum": fake0.double().sum().item(),
            "dreal0": dreal0.reshape(-1).numpy(),
            "D_losses": np.array(tr.D_losses, dtype=np.float64), "G_losses": np.array(tr.G_losses, dtype=np.float64),
        }
        for k, v in g0.items():
            out["g0." + k] = v.numpy()
        for k, v in d0.items():
            out["d0." + k] = v.numpy()
        for k, v in g1.items():
            out["g1." + k] = v.numpy()
        for k, v in d1.items():
            out["d1." + k] = v.numpy()
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes;", "D_losses", out["D_losses"], "G_losses", out["G_losses"],
          "keys", len(out))


if __name__ == "__main__":
    main()
