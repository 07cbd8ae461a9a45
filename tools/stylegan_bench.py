"""BASELINE config 4: StyleGAN2 128^3 (channel_base 4096, w_dim = z = 512, batch 8) on the B200
conv path: generator forward (the configuration BASELINE names) and the full train step
(iterD x StyleGAN2Loss.step_D + step_G at a step without the lazy R1 / path-length terms, Adam
included), plus - optionally - the v1-like generator `styleGAN_gen` (row a15).

Prints ms, voxels/s, the conv-only TFLOP/s with a per-signature conv table (CUDA events through
b200conv3d.functional.PROFILE), a per-kernel time table of one step (torch.profiler / CUPTI) and
the same work on stock torch layers under bf16 autocast (cuDNN witness: "what you get today").

  python tools/stylegan_bench.py --batch 8 --steps 3 [--witness] [--kernels] [--v1] [--json out.json]
"""
import argparse
import functools
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


def timed(fn, warmup, steps):
    for i in range(warmup):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def conv_table(fn):
    import b200conv3d.functional as BF
    BF.PROFILE = []
    fn(0)
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
    ms = sum(r["ms_total"] for r in rows)
    fl = sum(conv_flops(k) * r["n"] for k, r in agg.items())
    return {"conv_ms": ms, "conv_tflop": fl / 1e12, "conv_tflops": fl / max(ms, 1e-9) / 1e9, "rows": rows}


def kernel_table(fn, top=30):
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        fn(0)
        torch.cuda.synchronize()
    agg = {}
    for ev in prof.events():
        if ev.device_type is not None and "cuda" in str(ev.device_type).lower():
            r = agg.setdefault(ev.name[:90], [0.0, 0])
            r[0] += ev.device_time_total / 1e3 if hasattr(ev, "device_time_total") else ev.cuda_time_total / 1e3
            r[1] += 1
    rows = sorted(([k, v[0], v[1]] for k, v in agg.items()), key=lambda r: -r[1])
    return {"total_ms": sum(r[1] for r in rows), "rows": rows[:top]}


def op_table(fn, top=25):
    """ATen ops (not b2c launches) by device time, with input shapes: which glue op owns the generic kernels."""
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True) as prof:
        fn(0)
        torch.cuda.synchronize()
    tab = prof.key_averages(group_by_input_shape=True).table(sort_by="self_cuda_time_total", row_limit=top,
                                                             max_name_column_width=40, max_shapes_column_width=70)
    return tab


def build(kind, a, dev="cuda"):
    p = argparse.Namespace(z_size=a.z, filterG=a.cb, filterD=a.cb, device=dev, stylegan2=True, stylegan=False)
    torch.manual_seed(1234)
    if kind == "ours":
        import b200conv3d as B
        D, G = B.stylegan.Discriminator(p), B.stylegan.Generator(p)
        loss = B.stylegan.StyleGAN2Loss(p, G.mapping, G.synthesis, D, style_mixing_prob=0.9)
        G, D = G.to(dev), D.to(dev)
        return G, D, (lambda step, real, z: loss.step_D(step, real, z)), (lambda step, z: loss.step_G(step, z))
    import ref_stylegan as R
    D, G = R.Discriminator(a.cb).to(dev), R.Generator(a.z, a.cb).to(dev)
    pl = torch.zeros([], device=dev)
    return G, D, (lambda step, real, z: R.step_D(G, D, step, real, z)), (lambda step, z: R.step_G(G, D, step, z, pl))


def run(kind, a):
    dev = "cuda"
    G, D, step_D, step_G = build(kind, a)
    if kind == "ours":
        import b200conv3d.functional as BF
        BF.watch_optimizers()
    optD = torch.optim.Adam(D.parameters(), lr=1e-4, betas=(0.0, 0.9), fused=True)
    optG = torch.optim.Adam(G.parameters(), lr=5e-5, betas=(0.0, 0.9), fused=True)
    reals = [torch.rand(a.batch, 1, 128, 128, 128, device=dev) * 2 - 1 for _ in range(4)]
    gen = torch.Generator(device=dev).manual_seed(7)
    noise = lambda: torch.randn(a.batch, a.z, 1, 1, 1, device=dev, generator=gen)   # noqa: E731
    amp = functools.partial(torch.autocast, "cuda", dtype=torch.bfloat16)

    def fwd(i):
        with torch.no_grad(), amp():
            return G(noise())[0]

    def train(i):            # trainer.py:208-283 with the StyleGAN2 branches; step 1: no lazy regularisers
        for j in range(a.iterD):
            for q in D.parameters():
                q.requires_grad = True
            D.zero_grad()
            with amp():
                step_D(1, reals[(i * a.iterD + j) % 4], noise())
            optD.step()
            for q in D.parameters():
                q.requires_grad = False
        for q in G.parameters():
            q.requires_grad = True
        G.zero_grad()
        with amp():
            step_G(1, noise())
        optG.step()
        for q in G.parameters():
            q.requires_grad = False

    def fwd_bwd(i):          # "G fwd+bwd alone"
        for q in G.parameters():
            q.requires_grad = True
        G.zero_grad()
        with amp():
            img = G(noise())[0]
            img.float().square().mean().backward()
        for q in G.parameters():
            q.requires_grad = False

    for q in list(D.parameters()) + list(G.parameters()):
        q.requires_grad = False
    out = {"impl": kind}
    ms = timed(fwd, a.warmup, a.steps)
    out["gen_fwd"] = {"ms": ms, "voxels_per_sec": a.batch * VOX * 1e3 / ms}
    ms = timed(fwd_bwd, a.warmup, a.steps)
    out["gen_fwd_bwd"] = {"ms": ms, "voxels_per_sec": a.batch * VOX * 1e3 / ms}
    ms = timed(train, a.warmup, a.steps)
    out["train_step"] = {"ms": ms, "steps_per_sec": 1e3 / ms, "voxels_per_sec": a.batch * VOX * 1e3 / ms}
    out["peak_mem_gb"] = torch.cuda.max_memory_allocated() / 2 ** 30
    if kind == "ours":
        out["gen_fwd"].update(conv_table(fwd))
        out["train_step"].update(conv_table(train))
    if a.kernels:
        out["gen_fwd"]["kernels"] = kernel_table(fwd)
        out["train_step"]["kernels"] = kernel_table(train)
    if a.ops:
        print(op_table(train), file=sys.stderr)
    del G, D, optD, optG, reals
    torch.cuda.empty_cache()
    return out


def run_v1(kind, a):
    dev = "cuda"
    torch.manual_seed(4321)
    if kind == "ours":
        import b200conv3d as B
        G = B.stylegan.styleGAN_gen(argparse.Namespace(z_size=512)).to(dev)
    else:
        import ref_stylegan as R
        G = R.StyleGen1().to(dev)
    opt = torch.optim.Adam(G.parameters(), lr=5e-5, betas=(0.0, 0.9), fused=True)
    gen = torch.Generator(device=dev).manual_seed(7)

    def fwdbwd(i):
        G.zero_grad()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            y = G(torch.randn(a.batch, 512, 1, 1, 1, device=dev, generator=gen))
            y.float().square().mean().backward()
        opt.step()

    ms = timed(fwdbwd, a.warmup, a.steps)
    out = {"impl": kind, "v1_fwd_bwd": {"ms": ms, "voxels_per_sec": a.batch * VOX * 1e3 / ms}}
    if kind == "ours":
        out["v1_fwd_bwd"].update(conv_table(fwdbwd))
    if a.kernels:
        out["v1_fwd_bwd"]["kernels"] = kernel_table(fwdbwd)
    del G, opt
    torch.cuda.empty_cache()
    return out


def show(name, r):
    line = "%-8s %-10s %9.2f ms  %.3e voxels/s" % (r["impl"], name, r[name]["ms"], r[name]["voxels_per_sec"])
    if "conv_ms" in r[name]:
        line += "  conv %.1f ms (%.2f TF -> %.0f TFLOP/s)" % (r[name]["conv_ms"], r[name]["conv_tflop"], r[name]["conv_tflops"])
    print(line)
    for row in r[name].get("rows", [])[:24]:
        print("    %-6s T=%d %4d@%3d -> %4d@%3d k%d s%d calls %3d %8.2f ms %7.1f TF algo %d"
              % (row["pass"], row["transposed"], row["cin"], row["in"], row["cout"], row["out"], row["k"], row["stride"],
                 row["calls"], row["ms_total"], row["tflops"], row["algo"]))
    if "kernels" in r[name]:
        print("    kernels of one call: %.1f ms total" % r[name]["kernels"]["total_ms"])
        for k, ms, n in r[name]["kernels"]["rows"]:
            print("      %8.2f ms %5d  %s" % (ms, n, k))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--cb", type=int, default=4096, help="channel_base = filterG = filterD")
    ap.add_argument("--z", type=int, default=512)
    ap.add_argument("--iterD", type=int, default=2)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--witness", action="store_true", help="also time stock torch layers (cuDNN) under bf16 autocast")
    ap.add_argument("--ops", action="store_true", help="ATen op table of one train step (stderr)")
    ap.add_argument("--kernels", action="store_true", help="per-kernel time table of one call (torch.profiler)")
    ap.add_argument("--v1", action="store_true", help="also the v1-like generator styleGAN_gen")
    ap.add_argument("--json", default="")
    a = ap.parse_args()
    res = {"config": vars(a), "runs": []}
    t0 = time.time()
    for kind in ["ours"] + (["witness"] if a.witness else []):
        r = run(kind, a)
        res["runs"].append(r)
        show("gen_fwd", r), show("gen_fwd_bwd", r), show("train_step", r)
        print("%-8s peak mem %.1f GB" % (kind, r["peak_mem_gb"]))
        if a.v1:
            r1 = run_v1(kind, a)
            res["runs"].append(r1)
            show("v1_fwd_bwd", r1)
    print("wall %.0f s" % (time.time() - t0))
    if a.json:
        json.dump(res, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
