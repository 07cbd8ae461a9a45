"""BASELINE config 3: BigGAN-deep 128^3 train step (`--biggan=True --hinge=True --filterG=64
--filterD=64`, run.sh:1; batch 16, iterD 2, z 512 = main.py defaults) on the B200 conv path.

Prints ms/step, steps/s, voxels/s, the conv-only TFLOP/s and a per-signature conv table
(launches bracketed by CUDA events through b200conv3d.functional.PROFILE), plus the same
step on stock torch layers under bf16 autocast (cuDNN witness: "what you get today").

  python tools/biggan_bench.py --batch 16 --steps 3 [--witness] [--json out.json]
"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
VOX = 128 ** 3


def conv_flops(k):
    pas, tr, n, xc, xd, yc, yd, kk, stride = k[:9]
    small = xd if tr else yd
    return 2.0 * n * small ** 3 * xc * yc * kk ** 3


def run(kind, a):
    import b200conv3d as B
    import b200conv3d.functional as BF
    from b200conv3d.train import GanStep
    dev = "cuda"
    torch.manual_seed(1234)
    if kind == "ours":
        g, d = B.biggan.build(a.z, a.width, a.width, dev)
    else:
        import ref_biggan
        g, d = ref_biggan.Generator(a.z, a.width).to(dev), ref_biggan.Discriminator(a.width).to(dev)
        if a.channels_last:
            g, d = g.to(memory_format=torch.channels_last_3d), d.to(memory_format=torch.channels_last_3d)
    step = GanStep(g, d, a.z, hinge=True, iterD=a.iterD, autocast_dtype=torch.bfloat16, seed=99)
    reals = [torch.rand(a.batch, 1, 128, 128, 128, device=dev) * 2 - 1 for _ in range(4)]

    def one(i):
        step.step(reals[(i * a.iterD + j) % 4] for j in range(a.iterD))

    for i in range(a.warmup):
        one(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.steps):
        one(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    out = {"impl": kind, "ms_per_step": ms, "steps_per_sec": 1e3 / ms, "voxels_per_sec": a.batch * VOX * 1e3 / ms,
           "peak_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30}
    if kind == "ours":
        BF.PROFILE = []
        one(0)
        torch.cuda.synchronize()
        prof, BF.PROFILE = BF.PROFILE, None
        agg = {}
        for key, s0, s1, algo in prof:
            r = agg.setdefault(key, {"ms": 0.0, "n": 0, "algo": algo})
            r["ms"] += s0.elapsed_time(s1)
            r["n"] += 1
        rows = [{"pass": k[0], "transposed": k[1], "n": k[2], "cin": k[3], "in": k[4], "cout": k[5], "out": k[6],
                 "k": k[7], "stride": k[8], "calls": r["n"], "ms_total": r["ms"], "algo": r["algo"],
                 "tflops": conv_flops(k) * r["n"] / (r["ms"] * 1e-3) / 1e12} for k, r in agg.items()]
        rows.sort(key=lambda r: -r["ms_total"])
        conv_ms = sum(r["ms_total"] for r in rows)
        fl = sum(conv_flops(k) * r["n"] for k, r in agg.items())
        out.update({"conv_ms": conv_ms, "conv_tflop_per_step": fl / 1e12, "conv_tflops": fl / conv_ms / 1e9, "rows": rows})
    del step, g, d, reals
    torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--width", type=int, default=64)
    ap.add_argument("--z", type=int, default=512)
    ap.add_argument("--iterD", type=int, default=2)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--witness", action="store_true", help="also time stock torch layers (cuDNN) under bf16 autocast")
    ap.add_argument("--channels-last", action="store_true")
    ap.add_argument("--json", default="")
    a = ap.parse_args()
    res = {"config": vars(a), "runs": []}
    t0 = time.time()
    r = run("ours", a)
    res["runs"].append(r)
    print("ours   : %.1f ms/step  %.3f steps/s  %.3e voxels/s  conv %.1f ms (%.1f TF/step -> %.0f TFLOP/s)  peak mem %.1f GB"
          % (r["ms_per_step"], r["steps_per_sec"], r["voxels_per_sec"], r["conv_ms"], r["conv_tflop_per_step"],
             r["conv_tflops"], r["peak_mem_gb"]))
    for row in r["rows"][:28]:
        print("  %-6s T=%d %4d@%3d -> %4d@%3d k%d s%d calls %3d %8.2f ms %7.1f TF algo %d"
              % (row["pass"], row["transposed"], row["cin"], row["in"], row["cout"], row["out"], row["k"], row["stride"],
                 row["calls"], row["ms_total"], row["tflops"], row["algo"]))
    if a.witness:
        w = run("witness", a)
        res["runs"].append(w)
        print("witness (stock torch bf16 autocast, cuDNN): %.1f ms/step  %.3f steps/s  peak mem %.1f GB"
              % (w["ms_per_step"], w["steps_per_sec"], w["peak_mem_gb"]))
    print("wall %.0f s" % (time.time() - t0))
    if a.json:
        json.dump(res, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
