"""Generates tests/golden/dcgan_ref.npz by running the UNMODIFIED reference
(/root/reference: dcgan.py, trainer.py) on the CPU of the build container.

Only three modules that carry no arithmetic are stubbed so that the reference imports
(SURVEY.md F6/F7): `msl` (file missing from the reference), `pytorch_fid_wrapper`
(logging), `carbontracker.tracker` (all uses commented out).  Everything recorded here
comes out of the reference's own Generator / Discriminator / Trainer.D_step / G_step.

  PYTHONDONTWRITEBYTECODE=1 python oracle/gen_golden.py

The golden file pins oracle/ref_dcgan.py (tests/test_oracle_models.py), which in turn is
the CPU baseline of bench.py and the module-level oracle of the GPU tests.
"""
import argparse
import os
import sys
import tempfile
import types

import numpy as np
import torch

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "dcgan_ref.npz")


def stub_modules():
    msl = types.ModuleType("msl")
    msl.RandomCrop3D = object
    fid = types.ModuleType("pytorch_fid_wrapper")
    fid.set_config = lambda **k: None
    fid.fid = lambda *a, **k: 0.0
    ct = types.ModuleType("carbontracker")
    ctt = types.ModuleType("carbontracker.tracker")
    ctt.CarbonTracker = object
    ct.tracker = ctt
    sys.modules.update({"msl": msl, "pytorch_fid_wrapper": fid, "carbontracker": ct, "carbontracker.tracker": ctt})


def params(log_dir, width, z, batch):
    return argparse.Namespace(
        niters=1, batch_size=batch, z_size=z, filterG=width, filterD=width, iterD=1, lrG=5e-5, lrD=1e-4,
        data_path="", ngpu=1, steps_per_log=10 ** 9, steps_per_img_log=10 ** 9, log_dir=log_dir, device="cpu",
        load_params=False, dcgan=True, hybrid=False, stylegan2=False, stylegan=False, msl=False, sngan=False,
        sagan=False, biggan=False, hinge=False)


def main():
    stub_modules()
    sys.path.insert(0, REF)
    sys.dont_write_bytecode = True
    import trainer as ref_trainer          # noqa: E402  (the reference, unmodified)

    width, z, batch = 2, 8, 2
    torch.manual_seed(20261104)
    with tempfile.TemporaryDirectory() as tmp:
        tr = ref_trainer.Trainer([torch.zeros(1)] * 8, params(tmp, width, z, batch))
        g0 = {k: v.clone() for k, v in tr.netG.state_dict().items()}
        # D: keep only the conv filters; LayerNorm affine is the torch default (ones / zeros)
        d0 = {k: v.clone() for k, v in tr.netD.state_dict().items() if v.dim() == 5 and v.shape[2] == 4}
        torch.manual_seed(5)               # inputs are re-drawn from this seed by the tests
        zin = torch.randn(batch, z, 1, 1, 1)
        real = torch.rand(batch, 1, 128, 128, 128) * 2 - 1
        with torch.no_grad():
            fake0 = tr.netG(zin)
            dreal0 = tr.netD(real)
        # the two trainer phases, noise drawn by the reference from the global RNG
        for p in tr.netD.parameters():
            p.requires_grad = False
        for p in tr.netG.parameters():
            p.requires_grad = False
        torch.manual_seed(77)
        tr.D_step(0, real)
        torch.manual_seed(78)
        tr.G_step(0)
        g1 = tr.netG.state_dict()
        d1 = tr.netD.state_dict()
        out = {
            "torch_version": np.array(torch.__version__), "width": width, "z": z, "batch": batch,
            "zin": zin.numpy(), "real_seeded_by": np.array("torch.manual_seed(5); zin = randn(B,z,1,1,1); real = rand(B,1,128,128,128)*2-1"),
            "real_sample": real[:, 0, ::16, ::16, ::16].numpy(),
            "fake0_sample": fake0[:, 0, ::8, ::8, ::8].numpy(), "fake0_sum": fake0.double().sum().item(),
            "dreal0": dreal0.reshape(-1).numpy(),
            "D_losses": np.array(tr.D_losses, dtype=np.float64), "G_losses": np.array(tr.G_losses, dtype=np.float64),
        }
        for k, v in g0.items():
            out["g0." + k] = v.numpy()
        for k, v in d0.items():
            out["d0." + k] = v.numpy()
        # after one D_step + G_step: conv filters and a strided sample of the first LayerNorm affine
        for k, v in g1.items():
            if v.dim() == 5:
                out["g1." + k] = v.numpy()
        for k, v in d1.items():
            if v.dim() == 5 and v.shape[2] == 4:
                out["d1." + k] = v.numpy()
        out["d1.ln1_weight_sample"] = d1["main.1.weight"][:, ::16, ::16, ::16].numpy()
        out["d1.ln1_bias_sample"] = d1["main.1.bias"][:, ::16, ::16, ::16].numpy()
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes;", "D_losses", out["D_losses"], "G_losses", out["G_losses"])


if __name__ == "__main__":
    main()
