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


class Spy(TorchDispatchMode):
    def __torch_dispatch__(self, func, types, args=(), kwargs=None):
        name = str(func)
        if any(k in name for k in ("copy_", "clone", "_to_copy", "contiguous", "add", "mul")):
            big = [a for a in args if isinstance(a, torch.Tensor) and a.numel() >= (1 << 26)]
            if big:
                fr = [f for f in traceback.extract_stack()[:-1] if "b200conv3d" in f.filename or "stylegan" in f.filename]
                key = (name, tuple(big[0].shape), tuple(big[0].stride()), big[0].dtype,
                       " <- ".join("%s:%d" % (os.path.basename(f.filename), f.lineno) for f in fr[-5:]))
                SEEN[key] += 1
        return func(*args, **(kwargs or {}))


def main():
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
