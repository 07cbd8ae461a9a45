"""Who issues the large strided copies in one StyleGAN2 train step?  A TorchDispatchMode logs every aten copy /
clone / contiguous-like op on >= 2^26 elements with the Python stack that issued it (custom-Function backward frames
included)."""
import argparse
import collections
import os
import sys
import traceback

import torch
from torch.utils._python_dispatch import TorchDispatchMode

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import stylegan_bench as SB  # noqa: E402

SEEN = collections.Counter()
ANY = False
THRESH = 1 << 26


class Spy(TorchDispatchMode):
    def __torch_dispatch__(self, func, types, args=(), kwargs=None):
        name = str(func)
        views = ("permute", "view", "detach", "empty", "reshape", "alias", "as_strided", "expand", "slice", "select",
                 "transpose", "unsqueeze", "squeeze", "aten.t.", "unbind", "split", "narrow", "zeros", "ones")
        if (ANY and not any(v in name for v in views)) or any(k in name for k in ("copy_", "clone", "_to_copy",
                                                                                   "contiguous", "add", "mul")):
            big = [a for a in args if isinstance(a, torch.Tensor) and a.numel() >= THRESH]
            if big:
                fr = [f for f in traceback.extract_stack()[:-1] if "b200conv3d" in f.filename or "stylegan" in f.filename]
                key = (name, tuple(big[0].shape), tuple(big[0].stride()), big[0].dtype,
                       " <- ".join("%s:%d" % (os.path.basename(f.filename), f.lineno) for f in fr[-5:]))
                SEEN[key] += 1
        return func(*args, **(kwargs or {}))


def main_biggan():
    import b200conv3d as B
    from b200conv3d.train import GanStep
    torch.manual_seed(1234)
    g, d = B.biggan.build(512, 64, 64, "cuda")
    step = GanStep(g, d, 512, hinge=True, iterD=2, autocast_dtype=torch.bfloat16, seed=99)
    reals = [torch.rand(16, 1, 128, 128, 128, device="cuda") * 2 - 1 for _ in range(2)]
    for _ in range(2):
        step.step(iter(reals))
    torch.cuda.synchronize()
    with Spy():
        step.step(iter(reals))
    torch.cuda.synchronize()
    byb = collections.Counter()
    for k, n in SEEN.items():
        numel = 1
        for v in k[1]:
            numel *= v
        byb[k] = n * numel
    for k, b in sorted(byb.items(), key=lambda kv: -kv[1])[:40]:
        print("%6.2f Gelem  x%d  %s" % (b / 1e9, SEEN[k], k))


def main_dcgan():
    import b200conv3d as B
    from b200conv3d.train import GanStep
    torch.manual_seed(1234)
    g, d = B.dcgan.build(512, 128, 128, 128, "cuda")
    step = GanStep(g, d, 512, iterD=2, autocast_dtype=torch.bfloat16, seed=99)
    reals = [torch.rand(32, 1, 128, 128, 128, device="cuda") * 2 - 1 for _ in range(2)]
    for _ in range(2):
        step.step(iter(reals))
    torch.cuda.synchronize()
    global THRESH
    THRESH = 1 << 22
    with Spy():
        step.step(iter(reals))
    torch.cuda.synchronize()
    for k, n in sorted(SEEN.items(), key=lambda kv: -kv[1])[:40]:
        print("x%d  %s" % (n, k))


def main():
    global ANY
    if len(sys.argv) > 1 and sys.argv[1] == "dcgan":
        ANY = True
        return main_dcgan()
    if len(sys.argv) > 1 and sys.argv[1] == "biggan":
        ANY = True
        return main_biggan()
    a = argparse.Namespace(batch=8, cb=4096, z=512, iterD=2, steps=1, warmup=1, witness=False, kernels=False, v1=False,
                           json="", ops=False)
    orig = SB.timed

    def timed(fn, warmup, steps):
        ms = orig(fn, warmup, steps)
        if fn.__name__ == "train":
            with Spy():
                fn(0)
            torch.cuda.synchronize()
        return ms
    SB.timed = timed
    SB.run("ours", a)
    for k, n in sorted(SEEN.items(), key=lambda kv: -kv[1] * 1):
        print(n, k)


if __name__ == "__main__":
    main()
