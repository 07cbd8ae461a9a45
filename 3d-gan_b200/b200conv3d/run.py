"""Launcher: run an UNMODIFIED script of the reference on the B200 layers.

    python -m b200conv3d.run [--amp bf16|keep] [--no-fuse] main.py --dcgan=True --batch_size 32 ...
    torchrun --nproc-per-node 8 -m b200conv3d.run main.py --ngpu 8 --batch_size 4 ...

What it does before handing control to the script with ``runpy`` (the script and the reference's
trainer.py / dcgan.py / biggan.py / stylegan.py / FID_ResNet.py are executed byte for byte):

* :func:`b200conv3d.install` - ``nn.Conv3d / ConvTranspose3d / BatchNorm3d / LayerNorm / ReLU / LeakyReLU /
  Tanh / MaxPool3d / AvgPool3d`` and ``F.conv3d / F.conv_transpose3d / F.interpolate`` build / call the sm_100a
  library (the depthwise ``F.conv3d(groups=C)`` FIR of stylegan.py:197 included);
* ``--amp bf16`` (default): ``torch.cuda.amp.autocast`` - what trainer.py:14,213,259 imports - becomes a bf16
  autocast (BASELINE config 2 asks for bf16; the conv kernels compute in bf16 under any autocast, and with a
  bf16 autocast the few ATen ops around them - linears, attention - use the same 2-byte type).  The trainer's
  ``GradScaler`` keeps running: in bf16 its scaling is harmless.  ``--amp keep`` leaves torch's fp16 autocast;
* every ``nn.Sequential`` built afterwards gets its ``layer -> activation`` pairs fused (:func:`b200conv3d.fuse`)
  the first time it runs, unless ``--no-fuse``;
* under ``torchrun`` (WORLD_SIZE > 1): one process per GPU.  ``nn.DataParallel`` (trainer.py:70-72, reached with the
  reference's own ``--ngpu N`` flag) is rebound to :class:`ReplicaParallel`: a transparent wrapper (same
  ``module.`` state_dict prefix) whose gradients are all-reduced over NCCL by :class:`b200conv3d.parallel.GradReducer`
  - ``begin()`` rides on the trainer's ``zero_grad()`` (trainer.py:211,258), ``finish()`` on an optimizer pre-step
  hook - so every rank trains on its own ``--batch_size`` samples with per-replica BatchNorm statistics, exactly
  DataParallel's arithmetic.  Ranks > 0 do not write checkpoints or images.
"""
import argparse
import os
import runpy
import sys

import torch
import torch.nn as tnn


class _Bf16Autocast(torch.autocast):
    """Drop-in for ``torch.cuda.amp.autocast`` (same constructor) that always autocasts to bf16."""

    def __init__(self, enabled=True, dtype=torch.bfloat16, cache_enabled=True):
        super().__init__("cuda", enabled=enabled and torch.cuda.is_available(), dtype=torch.bfloat16,
                         cache_enabled=cache_enabled)


class ReplicaParallel(tnn.Module):
    """Stands in for ``nn.DataParallel(module)`` inside one rank of a torchrun job: forwards to ``module`` (the
    full replica of this rank) and averages its gradients over the ranks before the optimizer steps."""

    _by_param = {}            # id(param) -> reducer, for the optimizer pre-step hook
    _hooked = [False]

    def __init__(self, module, device_ids=None, output_device=None, dim=0):
        super().__init__()
        from .parallel import GradReducer
        self.module = module
        self.reducer = GradReducer(module)
        for p in module.parameters():
            ReplicaParallel._by_param[id(p)] = self.reducer
        if not ReplicaParallel._hooked[0]:
            from torch.optim.optimizer import register_optimizer_step_pre_hook
            register_optimizer_step_pre_hook(ReplicaParallel._before_step)
            ReplicaParallel._hooked[0] = True
        self._armed = False

    @staticmethod
    def _before_step(optimizer, *_a, **_k):
        seen = set()
        for g in optimizer.param_groups:
            for p in g["params"]:
                r = ReplicaParallel._by_param.get(id(p))
                if r is not None and id(r) not in seen:
                    seen.add(id(r))
                    r.finish()

    def zero_grad(self, set_to_none=True):
        # trainer.py:209-211 / 255-258: requires_grad flags are set, then zero_grad(), then backward
        super().zero_grad(set_to_none=set_to_none)
        self.reducer.begin()

    def forward(self, *a, **k):
        return self.module(*a, **k)


def _lazy_fuse_sequentials():
    """Fuse ``layer -> activation`` pairs of a Sequential right before its first forward (the reference applies its
    weight init after construction, dcgan.py:72,183, so construction time is too early to touch anything)."""
    from . import patch
    stock_forward = tnn.Sequential.forward

    def forward(self, *a, **k):
        if not getattr(self, "_b2c_fused", False):
            object.__setattr__(self, "_b2c_fused", True)
            patch.fuse(self)
        return stock_forward(self, *a, **k)

    tnn.Sequential.forward = forward


def setup(amp="bf16", fuse=True):
    """Everything the launcher does except running the script; returns (rank, world)."""
    import b200conv3d
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    if world > 1:
        import torch.distributed as dist
        local = int(os.environ.get("LOCAL_RANK", rank))
        if torch.cuda.is_available():
            torch.cuda.set_device(local)
        if not dist.is_initialized():
            dist.init_process_group("nccl" if torch.cuda.is_available() else "gloo")
        tnn.DataParallel = ReplicaParallel
        if rank > 0:
            torch.save = lambda *a, **k: None
            try:
                import torchvision.utils as vutils
                vutils.save_image = lambda *a, **k: None
            except Exception:
                pass
    b200conv3d.install()
    if amp == "bf16":
        import torch.cuda.amp as cuda_amp
        cuda_amp.autocast = _Bf16Autocast
    if fuse:
        _lazy_fuse_sequentials()
    return rank, world


def main(argv=None):
    ap = argparse.ArgumentParser(prog="python -m b200conv3d.run", description=__doc__.split("\n\n")[0])
    ap.add_argument("--amp", choices=("bf16", "keep"), default="bf16")
    ap.add_argument("--no-fuse", action="store_true")
    ap.add_argument("script")
    ap.add_argument("args", nargs=argparse.REMAINDER)
    ns = ap.parse_args(argv)
    setup(ns.amp, not ns.no_fuse)
    script = os.path.abspath(ns.script)
    sys.path.insert(0, os.path.dirname(script))
    sys.argv = [script] + ns.args
    runpy.run_path(script, run_name="__main__")


if __name__ == "__main__":
    main()
