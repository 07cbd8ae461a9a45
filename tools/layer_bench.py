"""Per-layer timing of the conv hot path through the C ABI (include/b2c.h).

Times fwd / dgrad / wgrad of each conv signature of a BASELINE config with CUDA
events on the launching stream and prints achieved TFLOP/s (or GB/s for the
HBM-bound degenerate layers) against MEASURED_PEAKS.json.  Not a bench line —
bench.py is the contract; this is the per-layer roofline table DESIGN.md cites.

  python tools/layer_bench.py --config dcgan --batch 32 [--iters 5] [--json out.json]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from abi import Abi, make_desc  # noqa: E402
from b200conv3d import _lib  # noqa: E402

# (name, transposed, cin, cout, k, stride, pad, dil, in_spatial)   reference site
CONFIGS = {
    # dcgan.py:45-70 (G) and dcgan.py:158-182 (D), filterG=filterD=128, z=512
    "dcgan": [
        ("G.L1 convT 512->2048 1->4", True, 512, 2048, 4, 1, 0, 1, 1),
        ("G.L2 convT 2048->1024 4->8", True, 2048, 1024, 4, 2, 1, 1, 4),
        ("G.L3 convT 1024->512 8->16", True, 1024, 512, 4, 2, 1, 1, 8),
        ("G.L4 convT 512->256 16->32", True, 512, 256, 4, 2, 1, 1, 16),
        ("G.L5 convT 256->128 32->64", True, 256, 128, 4, 2, 1, 1, 32),
        ("G.L6 convT 128->1 64->128", True, 128, 1, 4, 2, 1, 1, 64),
        ("D.L1 conv 1->128 128->64", False, 1, 128, 4, 2, 1, 1, 128),
        ("D.L2 conv 128->256 64->32", False, 128, 256, 4, 2, 1, 1, 64),
        ("D.L3 conv 256->512 32->16", False, 256, 512, 4, 2, 1, 1, 32),
        ("D.L4 conv 512->1024 16->8", False, 512, 1024, 4, 2, 1, 1, 16),
        ("D.L5 conv 1024->2048 8->4", False, 1024, 2048, 4, 2, 1, 1, 8),
        ("D.L6 conv 2048->1 4->1", False, 2048, 1, 4, 1, 0, 1, 4),
    ],
    # FID_ResNet.py:65-102 representative layers at batch 32
    "fid": [
        ("layer1 3x3x3 64->64 @32", False, 64, 64, 3, 1, 1, 1, 32),
        ("layer1 1x1 64->256 @32", False, 64, 256, 1, 1, 0, 1, 32),
        ("layer2 3x3x3 128->128 @16", False, 128, 128, 3, 1, 1, 1, 16),
        ("layer3 3x3x3 d2 256->256 @16", False, 256, 256, 3, 1, 2, 2, 16),
        ("layer4 3x3x3 d4 512->512 @16", False, 512, 512, 3, 1, 4, 4, 16),
        ("layer4 1x1 512->2048 @16", False, 512, 2048, 1, 1, 0, 1, 16),
        ("layer4 1x1 2048->512 @16", False, 2048, 512, 1, 1, 0, 1, 16),
    ],
    # utils.py:47-132 (BigGAN-deep hidden convs, f=64) at batch 16
    "biggan": [
        ("3x3x3 32->32 @128", False, 32, 32, 3, 1, 1, 1, 128),
        ("3x3x3 64->64 @64", False, 64, 64, 3, 1, 1, 1, 64),
        ("3x3x3 128->128 @32", False, 128, 128, 3, 1, 1, 1, 32),
        ("3x3x3 256->256 @16", False, 256, 256, 3, 1, 1, 1, 16),
        ("1x1 64->32 @128", False, 64, 32, 1, 1, 0, 1, 128),
        ("1x1 32->64 @128", False, 32, 64, 1, 1, 0, 1, 128),
        ("1x1 128->32 @64", False, 128, 32, 1, 1, 0, 1, 64),
        ("1x1 32->128 @64", False, 32, 128, 1, 1, 0, 1, 64),
        ("1x1 64->256 @32", False, 64, 256, 1, 1, 0, 1, 32),
        ("1x1 256->64 @32", False, 256, 64, 1, 1, 0, 1, 32),
        ("stem 3x3x3 1->64 @128", False, 1, 64, 3, 1, 1, 1, 128),
        ("head 3x3x3 64->1 @128", False, 64, 1, 3, 1, 1, 1, 128),
    ],
}


def out_sp(tr, v, k, s, p, d):
    return (v - 1) * s - 2 * p + d * (k - 1) + 1 if tr else (v + 2 * p - d * (k - 1) - 1) // s + 1


def time_it(fn, iters, flush):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.add_(1)                       # > L2: evict operands between timed launches
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
    ap.add_argument("--config", default="dcgan")
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--only", default="")
    ap.add_argument("--passes", default="fwd,dgrad,wgrad")
    ap.add_argument("--json", default="")
    ap.add_argument("--bias", action="store_true", help="forward with a fused bias + ReLU epilogue (utils.py:8-14 convs)")
    a = ap.parse_args()
    peaks = {}
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peaks = json.load(open(pk))
    tf_peak = peaks.get("bf16_tflops", 1590.0)
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    gpu = Abi(_lib.PRODUCT_LIB, "cuda:0")
    flush = torch.zeros(256 << 20, dtype=torch.uint8, device="cuda")   # 256 MB > 126 MB L2
    bf = torch.bfloat16
    rows = []
    print("%-34s %-6s %9s %10s %8s  %s" % ("layer", "pass", "ms", "TFLOP/s", "%peak", "algo"))
    for name, tr, cin, cout, k, s, p, d, sp in CONFIGS[a.config]:
        if a.only and a.only not in name:
            continue
        n = a.batch
        osp = out_sp(tr, sp, k, s, p, d)
        desc = make_desc(tr, (n, cin, sp, sp, sp), (n, cout, osp, osp, osp), (k,) * 3, (s,) * 3, (p,) * 3,
                         (d,) * 3, bf, 0)
        x = torch.randn(n, sp, sp, sp, cin, device="cuda").to(bf)
        dy = torch.randn(n, osp, osp, osp, cout, device="cuda").to(bf)
        w = torch.randn((cin, cout, k, k, k) if tr else (cout, cin, k, k, k), device="cuda") * 0.02
        small_vox = n * (sp if tr else osp) ** 3
        flops = 2.0 * small_vox * cin * cout * k ** 3
        io_bytes = 2.0 * (x.numel() + dy.numel()) + 2.0 * w.numel()
        import ctypes as C
        wp_f = gpu.pack(desc, 0, w)
        wp_d = gpu.pack(desc, 1, w)
        y = torch.empty_like(dy)
        dx = torch.empty_like(x)
        dw = torch.empty_like(w)
        wsn = max(gpu.lib.b2c_conv3d_workspace(C.byref(desc), i) for i in range(3))
        ws = torch.empty(max(wsn, 16), dtype=torch.uint8, device="cuda")
        st = gpu._st()
        P = gpu._p

        from b200conv3d._lib import Fusion
        bias = torch.randn(cout, device="cuda")
        fu = C.byref(Fusion(P(bias), 1, 0.0)) if a.bias else None

        def fwd():
            gpu._ck(gpu.lib.b2c_conv3d_fwd(C.byref(desc), P(x), P(wp_f), P(y), fu, P(ws), ws.numel(), st), "fwd")

        def dgrad():
            gpu._ck(gpu.lib.b2c_conv3d_dgrad(C.byref(desc), P(dy), P(wp_d), P(dx), P(ws), ws.numel(), st), "dgrad")

        def wgrad():
            gpu._ck(gpu.lib.b2c_conv3d_wgrad(C.byref(desc), P(x), P(dy), P(dw), None, P(ws), ws.numel(), st),
                    "wgrad")

        for pname, fn in (("fwd", fwd), ("dgrad", dgrad), ("wgrad", wgrad)):
            if pname not in a.passes.split(","):
                continue
            ms = time_it(fn, a.iters, flush)
            algo = gpu.last_algo()
            tfs = flops / ms / 1e9
            gbs = io_bytes / ms / 1e6
            rows.append(dict(layer=name, **{"pass": pname}, ms=ms, tflops=tfs, frac_tensor=tfs / tf_peak,
                             gbs=gbs, frac_hbm=gbs / hbm_peak, algo=algo, gflop=flops / 1e9))
            print("%-34s %-6s %9.3f %10.1f %7.1f%%  %s  (%.0f GB/s, %.1f%% hbm)" %
                  (name, pname, ms, tfs, 100 * tfs / tf_peak, {1: "direct", 2: "tc"}.get(algo, algo), gbs,
                   100 * gbs / hbm_peak))
        del x, dy, w, y, dx, dw, ws
    tot = sum(r["ms"] for r in rows)
    totf = sum(r["gflop"] for r in rows) / 1e3
    print("total %.2f ms for %.1f TFLOP -> %.1f TFLOP/s (%.1f%% of %.0f)" %
          (tot, totf, totf / tot * 1e3, 100 * totf / tot * 1e3 / tf_peak, tf_peak))
    if a.json:
        json.dump(dict(config=a.config, batch=a.batch, peaks=peaks, rows=rows), open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
