#!/usr/bin/env python
"""Narrow stride-1 3^3 convolutions (utils.py:57-60,105-109 hidden convs of BigGAN-deep, stylegan.py:285-287) through
the Python layer, where the lowering is chosen: W-window (b2c_window_w + block-Toeplitz filter, one tap along W)
against the older three-tap W-folding.  Times forward, and forward + backward (dx, dw), with CUDA events; reports
USEFUL TFLOP/s (2 * N * vox * Cin * Cout * 27 per pass).

  python tools/narrow_bench.py [--json out.json]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))

CASES = [  # name, batch, cin, cout, spatial
    ("biggan hidden 32->32 @128^3 b16", 16, 32, 32, 128),
    ("biggan hidden 64->64 @64^3 b16", 16, 64, 64, 64),
    ("biggan hidden 32->32 @64^3 b16", 16, 32, 32, 64),
    ("stylegan2 32->32 @128^3 b8", 8, 32, 32, 128),
    ("stylegan2 64->64 @64^3 b8", 8, 64, 64, 64),
]


def timed(fn, iters=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--json", default="")
    a = ap.parse_args()
    import b200conv3d.functional as BF
    rows = []
    for name, n, cin, cout, sp in CASES:
        x = torch.randn(n, sp, sp, sp, cin, device="cuda").to(torch.bfloat16).permute(0, 4, 1, 2, 3)
        w = (torch.randn(cout, cin, 3, 3, 3, device="cuda") * 0.05).requires_grad_(True)
        b = torch.zeros(cout, device="cuda", requires_grad=True)
        gy = torch.randn(n, sp, sp, sp, cout, device="cuda").to(torch.bfloat16).permute(0, 4, 1, 2, 3)
        flops = 2.0 * n * sp ** 3 * cin * cout * 27
        for mode in ("window", "fold"):
            BF.WINDOW[0] = mode == "window"

            def fwd():
                with torch.no_grad():
                    BF.conv3d(x, w, b, 1, 1)

            def fwd_bwd():
                xr = x.detach().requires_grad_(True)
                w.grad = b.grad = None
                BF.conv3d(xr, w, b, 1, 1).backward(gy)

            tf, tfb = timed(fwd), timed(fwd_bwd)
            r = {"layer": name, "lowering": mode, "fwd_ms": tf, "fwd_tflops_useful": flops / tf / 1e9,
                 "fwd_bwd_ms": tfb, "fwd_bwd_tflops_useful": 3 * flops / tfb / 1e9}
            rows.append(r)
            print("%-36s %-6s fwd %7.3f ms %7.1f TF/s   fwd+bwd %7.3f ms %7.1f TF/s" %
                  (name, mode, tf, r["fwd_tflops_useful"], tfb, r["fwd_bwd_tflops_useful"]))
        BF.WINDOW[0] = True
        del x, gy
        torch.cuda.empty_cache()
    if a.json:
        json.dump(rows, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
