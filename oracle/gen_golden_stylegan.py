"""Generates tests/golden/stylegan_ops_ref.npz from the UNMODIFIED reference functions
`stylegan.modulated_conv3d` / `stylegan.conv3d_resample` / `stylegan._upfirdn3d_ref`
(/root/reference/stylegan.py) for the call signatures its layers issue, and checks
oracle/ref_stylegan_ops.py against them in this container (forward, and the gradients w.r.t.
input, weight and styles).

  PYTHONDONTWRITEBYTECODE=1 python oracle/gen_golden_stylegan.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ref_stylegan_ops as O  # noqa: E402
from gen_golden import REF, stub_modules  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "stylegan_ops_ref.npz")


def grads(fn, x, w, s, nz, gy, kind):
    x, w, s = x.clone().requires_grad_(True), w.clone().requires_grad_(True), s.clone().requires_grad_(True)
    y = fn(x, w, s, nz)
    y.backward(gy)
    return y.detach(), x.grad, w.grad, (s.grad if kind == "mod" else None)


def main():
    stub_modules()
    sys.path.insert(0, REF)
    sys.dont_write_bytecode = True
    import stylegan as S          # noqa: E402  (the reference, unmodified)
    out = {"torch_version": np.array(torch.__version__)}
    for case in O.CASES:
        name, kind = case[0], case[1]
        x, w, s, nz, gy = O.case_inputs(*case)
        ref = grads(lambda a, b, c, d: O.run_case(S, case, a, b, c, d), x, w, s, nz, gy, kind)
        ora = grads(lambda a, b, c, d: O.run_case(O, case, a, b, c, d), x, w, s, nz, gy, kind)
        for r, o, what in zip(ref, ora, ("y", "dx", "dw", "ds")):
            if r is None:
                continue
            torch.testing.assert_close(o, r, rtol=1e-5, atol=1e-6, msg=lambda m: "%s %s: %s" % (name, what, m))
            out["%s.%s" % (name, what)] = r.numpy()
        print("%-16s y %s ok" % (name, tuple(ref[0].shape)))
    # the FIR funnel on its own, incl. the 1-channel image up-sampling of SynthesisBlock.forward (:620-634)
    f = O.make_filter()
    g = torch.Generator().manual_seed(99)
    for name, shape, kw in (("fir_img_up2", (2, 1, 4, 4, 4), dict(up=2, padding=[2, 1, 2, 1, 2, 1], gain=8)),
                            ("fir_down2", (2, 8, 8, 8, 8), dict(down=2, padding=[1, 1, 1, 1, 1, 1])),
                            ("fir_pad", (2, 8, 7, 7, 7), dict(padding=[2, 1, 2, 1, 2, 1])),
                            ("fir_crop", (2, 8, 9, 9, 9), dict(padding=[-1, -1, -1, -1, -1, -1]))):
        x = torch.randn(*shape, generator=g)
        y = S._upfirdn3d_ref(x, f, **kw)
        p = kw.get("padding")
        yo = O.upfirdn3d(x, f, up=kw.get("up", 1), down=kw.get("down", 1), pad=(p[0], p[1]), gain=kw.get("gain", 1))
        torch.testing.assert_close(yo, y, rtol=1e-6, atol=1e-6)
        out[name + ".x"], out[name + ".y"] = x.numpy(), y.numpy()
        print("%-16s y %s ok" % (name, tuple(y.shape)))
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
