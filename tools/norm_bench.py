"""Normalisation-pass roofline table (BN / LN passes are HBM bound).

Times the C-ABI LayerNorm(+LeakyReLU) and BatchNorm3d(+ReLU) passes at the DCGAN
128^3 shapes (dcgan.py:48-64 BN after G.L1..L5, dcgan.py:162-178 LN after D.L1..L5) and
prints achieved GB/s of ALGORITHMIC bytes:
  LN fwd  : 2 (stats read) + 2 (read) + 2 (write) bytes / element  + 8 |gamma| once
  LN bwd  : 4 (stats: x, dy) + 4 (x, dy) + 2 (dx)                   + 16 |gamma| (read + grads)
  BN fwd  : 2 + 2 + 2,   BN bwd : 4 + 4 + 2
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))


def timeit(fn, iters, flush):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--width", type=int, default=128)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--json", default="")
    a = ap.parse_args()
    import b200conv3d.functional as BF
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    hbm = float(peaks.get("hbm_gbs", 6540.0))
    dev = "cuda"
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rows = []
    N, f = a.batch, a.width
    layers = [("LN D.L%d" % (i + 1), "ln", f << i, 64 >> i) for i in range(5)] + \
             [("BN G.L%d" % (i + 1), "bn", (f * 16) >> i, 4 << i) for i in range(5)]
    print("%-10s %-4s %10s %9s %9s %7s" % ("layer", "pass", "elements", "ms", "GB/s", "%hbm"))
    for name, kind, c, s in layers:
        x = (torch.randn(N, c, s, s, s, device=dev, dtype=torch.bfloat16)).contiguous(memory_format=torch.channels_last_3d)
        x.requires_grad_(True)
        g = torch.randn_like(x)
        el = x.numel()
        if kind == "ln":
            ga = torch.ones(c, s, s, s, device=dev, requires_grad=True)
            be = torch.zeros(c, s, s, s, device=dev, requires_grad=True)
            fwd = lambda: BF.layer_norm_act(x, ga, be, 1e-5, "lrelu", 0.2)
            bytes_f, bytes_b = 6 * el + 8 * ga.numel(), 10 * el + 16 * ga.numel()
        else:
            ga = torch.ones(c, device=dev, requires_grad=True)
            be = torch.zeros(c, device=dev, requires_grad=True)
            rm, rv = torch.zeros(c, device=dev), torch.ones(c, device=dev)
            fwd = lambda: BF.batch_norm_act(x, ga, be, rm, rv, True, 0.1, 1e-5, "relu", 0.0)
            bytes_f, bytes_b = 6 * el, 10 * el
        tf = timeit(lambda: fwd(), a.iters, flush)
        y = fwd()

        def bwd():
            x.grad = ga.grad = be.grad = None
            y.backward(g, retain_graph=True)
        tb = timeit(bwd, a.iters, flush)
        for pas, t, b in (("fwd", tf, bytes_f), ("bwd", tb, bytes_b)):
            gbs = b / t / 1e6
            print("%-10s %-4s %10d %9.3f %9.0f %6.1f%%" % (name, pas, el, t, gbs, 100 * gbs / hbm))
            rows.append({"layer": name, "pass": pas, "elements": el, "ms": t, "algorithmic_bytes": b, "gbs": gbs, "frac_hbm": gbs / hbm})
        del x, g, y
    if a.json:
        json.dump({"batch": N, "hbm_peak_gbs": hbm, "rows": rows}, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
