"""Depthwise FIR resampling roofline point (stylegan.py:172-200 at StyleGAN2 config 4 sizes):
[8, 32, 131^3] -> [8, 32, 128^3] (the largest FIR after an up-sampling conv) and the stride-2
discriminator pre-filter [8, 32, 132^3] -> [8, 32, 129^3].  Algorithmic bytes = 2 (|x| + |y|)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))
import b200conv3d.functional as BF  # noqa: E402
import b200conv3d.stylegan_ops as P  # noqa: E402

f = P.make_resample_filter()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
hbm = 6540.0
pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
if os.path.exists(pk):
    hbm = float(json.load(open(pk)).get("hbm_gbs", hbm))
for name, shape, kw in (("G up-conv FIR 32ch 131^3->128^3", (8, 32, 131, 131, 131), dict(padding=[0, 0, 0, 0, 0, 0], gain=4)),
                        ("D pre-filter 32ch 128^3->129^3", (8, 32, 128, 128, 128), dict(padding=[2, 2, 2, 2, 2, 2])),
                        ("D skip FIR down2 32ch 128^3->64^3", (8, 32, 128, 128, 128), dict(down=2, padding=[1, 1, 1, 1, 1, 1])),
                        ("G FIR 64ch 67^3->64^3", (8, 64, 67, 67, 67), dict(padding=0, gain=4))):
    x = torch.randn(*shape, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last_3d).requires_grad_(True)
    y = BF.upfirdn3d(x, f, **kw)
    g = torch.randn_like(y)
    for pas, fn in (("fwd", lambda: BF.upfirdn3d(x, f, **kw)), ("bwd", lambda: y.backward(g, retain_graph=True))):
        fn()
        ts = []
        for _ in range(5):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        t = sorted(ts)[2]
        b = 2.0 * (x.numel() + y.numel())
        print("%-36s %-4s %8.3f ms  %7.0f GB/s  %5.1f%% hbm" % (name, pas, t, b / t / 1e6, 100 * b / t / 1e6 / hbm))
