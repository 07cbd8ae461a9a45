"""Per-kernel device-time table (torch.profiler / CUPTI) and per-signature conv table (CUDA events) of one call
of a secondary workload: `fid` (FID ResNet-50 features, batch 25) or `biggan` (one BigGAN-deep train step).

  python tools/kernel_table.py fid|biggan [--top 40]
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
from stylegan_bench import conv_table, kernel_table  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("workload", choices=["fid", "biggan"])
    ap.add_argument("--top", type=int, default=40)
    a = ap.parse_args()
    import b200conv3d as B
    dev = "cuda"
    torch.manual_seed(0)
    if a.workload == "fid":
        model = B.fid_resnet.resnet50().to(dev)
        x = torch.randn(25, 1, 128, 128, 128, device=dev)

        def fn(i):
            with torch.autocast("cuda", dtype=torch.bfloat16):
                return B.fid_resnet.features(model, x)
    else:
        from b200conv3d.train import GanStep
        g, d = B.biggan.build(512, 64, 64, dev)
        step = GanStep(g, d, 512, hinge=True, iterD=2, autocast_dtype=torch.bfloat16, seed=99)
        reals = [torch.rand(16, 1, 128, 128, 128, device=dev) * 2 - 1 for _ in range(2)]

        def fn(i):
            step.step(reals[j % 2] for j in range(2))
    for i in range(2):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn(0)
    e1.record()
    torch.cuda.synchronize()
    print("%s: %.2f ms per call" % (a.workload, e0.elapsed_time(e1)))
    ct = conv_table(fn)
    print("conv launches: %.2f ms, %.2f TF -> %.0f TFLOP/s" % (ct["conv_ms"], ct["conv_tflop"], ct["conv_tflops"]))
    for row in ct["rows"][:a.top]:
        print("  %-6s T=%d %4d@%3d -> %4d@%3d k%d s%d calls %3d %8.2f ms %7.1f TF algo %d"
              % (row["pass"], row["transposed"], row["cin"], row["in"], row["cout"], row["out"], row["k"], row["stride"],
                 row["calls"], row["ms_total"], row["tflops"], row["algo"]))
    kt = kernel_table(fn, top=a.top)
    print("kernels of one call: %.1f ms total" % kt["total_ms"])
    for k, ms, n in kt["rows"]:
        print("  %8.2f ms %5d  %s" % (ms, n, k))


if __name__ == "__main__":
    main()
