"""Loss-curve parity over N training steps (north_star: "matching loss curves over 100 steps";
SURVEY.md §8(c) G3).

Runs the DCGAN step schedule of trainer.py:208-303 twice on the same GPU with identical
initial weights, real batches and noise draws:
  ref  : stock torch layers (oracle/ref_dcgan.py models), fp32, TF32 off
  ours : b200conv3d drop-in layers, bf16 autocast (the product path)
and reports, per curve (D(x), D(G(z)), G loss), the mean |diff| relative to the range of
the reference curve, next to the same figure for stock torch layers under bf16 autocast
(the bf16 noise floor).  Criterion: < 5 % of the range, or no further from the
fp32 curve than 1.5x the stock bf16 run, and no NaN.
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def run(kind, a, init):
    import b200conv3d as B
    import ref_dcgan
    dev = "cuda"
    if kind in ("ref", "witness"):
        g, d = ref_dcgan.Generator(a.z, a.width).to(dev), ref_dcgan.Discriminator(a.width).to(dev)
        amp = None if kind == "ref" else torch.bfloat16
    else:
        g, d = B.dcgan.build(a.z, a.width, a.width, 128, "cpu", fuse=False)
        g, d = g.to(dev), d.to(dev)
        amp = torch.bfloat16
    g.load_state_dict(init[0])
    d.load_state_dict(init[1])
    if kind == "ours":
        B.fuse(g), B.fuse(d)
    step = B.train.GanStep(g, d, a.z, iterD=a.iterD, autocast_dtype=amp, seed=4242)
    data = torch.Generator(device=dev)
    data.manual_seed(777)
    curves = {"D_real": [], "D_fake": [], "G": []}
    for i in range(a.steps):
        for _ in range(a.iterD):
            real = torch.rand(a.batch, 1, 128, 128, 128, device=dev, generator=data) * 2 - 1
            er, ef = step.D_step(real)
        eg, _ = step.G_step(a.batch)
        curves["D_real"].append(er.item()), curves["D_fake"].append(ef.item()), curves["G"].append(eg.item())
    return curves


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--width", type=int, default=16)
    ap.add_argument("--z", type=int, default=64)
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--iterD", type=int, default=2)
    ap.add_argument("--json", default="")
    a = ap.parse_args()
    import ref_dcgan
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(0)
    init = (ref_dcgan.Generator(a.z, a.width).state_dict(), ref_dcgan.Discriminator(a.width).state_dict())
    ref = run("ref", a, init)
    ours = run("ours", a, init)
    wit = run("witness", a, init)      # stock torch layers under bf16 autocast (cuDNN): the bf16 noise floor
    out = {"steps": a.steps, "width": a.width, "batch": a.batch, "iterD": a.iterD, "curves": {}}
    ok = True
    for k in ref:
        r, o = torch.tensor(ref[k]), torch.tensor(ours[k])
        rng = (r.max() - r.min()).item()
        mad = (r - o).abs().mean().item()
        last = (r[-10:] - o[-10:]).abs().mean().item()
        rel = mad / max(rng, 1e-12)
        w = torch.tensor(wit[k])
        mad_w = (r - w).abs().mean().item()
        early = (r[:30] - o[:30]).abs().mean().item() / max((r[:30].max() - r[:30].min()).item(), 1e-12)
        out["curves"][k] = {"range": rng, "mean_abs_diff": mad, "rel": rel, "last10_abs_diff": last, "witness_mean_abs_diff": mad_w,
                            "first30_rel": early,
                            "ref_first_last": [ref[k][0], ref[k][-1]], "ours_first_last": [ours[k][0], ours[k][-1]]}
        print("%-7s range %.4f  mean|diff| %.5f (%.2f%% of range)  last-10 |diff| %.5f   ref %.4f -> %.4f   ours %.4f -> %.4f"
              % (k, rng, mad, 100 * rel, last, ref[k][0], ref[k][-1], ours[k][0], ours[k][-1]))
        print("        stock-torch bf16 autocast witness mean|diff| %.5f (%.2f%% of range); ours first-30-step rel %.2f%%"
              % (mad_w, 100 * mad_w / max(rng, 1e-12), 100 * early))
        # pass: within 5 % of the curve range, or no further from the fp32 curve than 1.5x what
        # stock PyTorch bf16 autocast is (an un-regularised WGAN amplifies any rounding difference)
        ok = ok and (rel < 0.05 or mad < 1.5 * mad_w) and all(map(lambda v: v == v, ours[k]))
    out["pass"] = bool(ok)
    out["ref"], out["ours"], out["witness"] = ref, ours, wit
    print("PASS" if ok else "FAIL")
    if a.json:
        json.dump(out, open(a.json, "w"), indent=1)
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
