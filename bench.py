#!/usr/bin/env python
"""bench.py — the driver's measurement contract for the 3D-GAN conv3d hot path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload dcgan|biggan|stylegan2|fid]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workloads (BASELINE.json configs; the default invocation is configs[1], the one the metric is quoted on):
  dcgan     configs[1]  3D DCGAN 128^3, batch 32 per GPU, bf16; step = iterD(2) x D_step + G_step (trainer.py:296-303)
  biggan    configs[2]  BigGAN-deep 128^3 (`--biggan --hinge --filterG=64 --filterD=64`, run.sh:1), batch 16 per GPU
  stylegan2 configs[3]  StyleGAN2 128^3 (channel_base 4096, z = w = 512), batch 8 per GPU, 2 x step_D + step_G
  fid       configs[4]  FID_ResNet.resnet50 features (eval_utils.fid_3d), 25 volumes per step per GPU, the [n,2048]
                        features all-gathered inside the timed region (10 000 volumes = --gpus 8 --steps 50)
Data-parallel over N GPUs (one process per GPU, NCCL gradient all-reduce overlapped with backward for the GAN
steps; independent shards + one all_gather for fid).  Synthetic data, random-init weights.

Prints ONE JSON line (rank 0).  `value` = units per second (generated voxels, or volumes for fid) with the inputs
already resident in HBM; `e2e` = the same step fed from pinned HOST memory through the public API with the step's
results read back; `roofline` = the tcgen05 conv launches of one profiled step (CUDA events on the launching stream)
against MEASURED_PEAKS.json; `witness` = the same step on stock torch layers under bf16 autocast (cuDNN) on the same
GPU; `cpu_baseline` / `--impl reference` = the reference's CPU path (stock torch layers on the host cores).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))

VOX = 128 ** 3
DEFAULT_BATCH = {"dcgan": 32, "biggan": 16, "stylegan2": 8, "fid": 25}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="dcgan", choices=sorted(DEFAULT_BATCH))
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (default: the BASELINE config's)")
    ap.add_argument("--global-batch", type=int, default=0,
                    help="strong scaling: split this global batch over the ranks (what --batch_size B --ngpu N means "
                         "in the reference, trainer.py:70-72,84)")
    ap.add_argument("--iterD", type=int, default=2)
    ap.add_argument("--width", type=int, default=0, help="filterG = filterD (dcgan 128, biggan 64, stylegan2 channel_base 4096)")
    ap.add_argument("--z", type=int, default=512)
    ap.add_argument("--two-pass-d", action="store_true",
                    help="dcgan: D(real), D(fake) as two passes, the reference's own schedule (trainer.py:239-241)")
    ap.add_argument("--fused-stats", default="all", choices=["all", "none", "channel", "sample"],
                    help="which norm statistics ride in the conv epilogue (measurement switch; default all)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--witness", dest="witness", action="store_true", default=None,
                    help="time stock torch layers under bf16 autocast (cuDNN) on the same GPU (default: on for dcgan at N=1)")
    ap.add_argument("--no-witness", dest="witness", action="store_false")
    ap.add_argument("--cpu-batch", type=int, default=0)
    ap.add_argument("--cpu-budget-s", type=float, default=150.0)
    ap.add_argument("--profile-json", default="", help="write the per-launch conv timing table here")
    a = ap.parse_args()
    a.width = a.width or {"dcgan": 128, "biggan": 64, "stylegan2": 4096, "fid": 0}[a.workload]
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if a.global_batch:
        if a.global_batch % world:
            raise SystemExit("--global-batch must be a multiple of the number of ranks")
        a.batch = a.global_batch // world
    a.batch = a.batch or DEFAULT_BATCH[a.workload]
    a.cpu_batch = a.cpu_batch or {"dcgan": 2, "biggan": 2, "stylegan2": 2, "fid": 2}[a.workload]
    return a


# --------------------------------------------------------------------------- clocks
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU every 100 ms through NVML."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons, self.max_mhz = index, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
                 nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake"}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.1)

    def result(self):
        self.stop_flag = True
        if self.nv is None or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def physical_cores():
    try:
        import psutil
        return psutil.cpu_count(logical=False) or os.cpu_count()
    except Exception:
        return os.cpu_count()


# ------------------------------------------------------------------ workload descriptions
def metric_of(a):
    if a.workload == "fid":
        return "fid_resnet50_3d_volumes_per_sec", "volumes/s"
    return {"dcgan": "dcgan128_train_voxels_per_sec", "biggan": "biggan128_train_voxels_per_sec",
            "stylegan2": "stylegan2_128_train_voxels_per_sec"}[a.workload], "voxels/s"


def units_per_step(a, batch):
    return batch if a.workload == "fid" else batch * VOX


def workload_text(a):
    return {
        "dcgan": "3D DCGAN 128^3 1-channel, train step = %d x D_step + G_step (trainer.py:296-303), BASELINE configs[1]" % a.iterD,
        "biggan": "3D BigGAN-deep 128^3 (--biggan --hinge, run.sh:1), train step = %d x D_step + G_step, BASELINE configs[2]" % a.iterD,
        "stylegan2": "3D StyleGAN2 128^3 (channel_base %d, modulated conv3d), train step = %d x step_D + step_G at a step "
                     "without the lazy R1 / path-length terms (stylegan.py:39-99), BASELINE configs[3]" % (a.width, a.iterD),
        "fid": "FID_ResNet.resnet50 features of synthetic 128^3 volumes (eval_utils.py:41-52, eval_metrics.py:72-73), "
               "BatchNorm in train mode, features all-gathered, BASELINE configs[4]",
    }[a.workload]


# ------------------------------------------------------------------ CPU reference arm (oracle/: allowed here only)
def cpu_reference(a, n_timed, n_warm, budget_s):
    """The reference's own CPU path (stock torch layers; oracle/ref_*.py restate dcgan.py / biggan.py + utils.py /
    stylegan.py / FID_ResNet.py + trainer.py and are pinned against the imported reference by tests/golden) on the host
    cores, fp32.  A bounded sample: dcgan runs whole train steps at batch `cpu_batch`; biggan / stylegan2 time ONE D
    step and ONE G step (train step = iterD x D + G); fid times forward batches."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    cores = physical_cores()
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    b = a.cpu_batch
    if a.workload == "fid":
        import ref_fid_resnet
        m = ref_fid_resnet.resnet50()

        def one():
            with torch.no_grad():
                m(torch.randn(b, 1, 128, 128, 128)).mean(dim=(2, 3, 4))
        composed = None
    elif a.workload == "dcgan":
        import ref_dcgan
        tr = ref_dcgan.RefTrainer(ref_dcgan.Generator(a.z, a.width), ref_dcgan.Discriminator(a.width), a.z)

        def one():
            for _ in range(a.iterD):
                tr.d_step(torch.rand(b, 1, 128, 128, 128) * 2 - 1)
            tr.g_step(b)
        composed = None
    elif a.workload == "biggan":
        import ref_biggan
        import ref_dcgan
        tr = ref_dcgan.RefTrainer(ref_biggan.Generator(a.z, a.width), ref_biggan.Discriminator(a.width), a.z, hinge=True)
        composed = (lambda: tr.d_step(torch.rand(b, 1, 128, 128, 128) * 2 - 1), lambda: tr.g_step(b))
    else:
        import ref_stylegan as R
        D, G = R.Discriminator(a.width), R.Generator(a.z, a.width)
        optD = torch.optim.Adam(D.parameters(), lr=1e-4, betas=(0.0, 0.9))
        optG = torch.optim.Adam(G.parameters(), lr=5e-5, betas=(0.0, 0.9))
        pl = torch.zeros([])
        for q in list(D.parameters()) + list(G.parameters()):
            q.requires_grad = False

        def d_phase():
            for q in D.parameters():
                q.requires_grad = True
            D.zero_grad()
            R.step_D(G, D, 1, torch.rand(b, 1, 128, 128, 128) * 2 - 1, torch.randn(b, a.z, 1, 1, 1))
            optD.step()
            for q in D.parameters():
                q.requires_grad = False

        def g_phase():
            for q in G.parameters():
                q.requires_grad = True
            G.zero_grad()
            R.step_G(G, D, 1, torch.randn(b, a.z, 1, 1, 1), pl)
            optG.step()
            for q in G.parameters():
                q.requires_grad = False
        composed = (d_phase, g_phase)

    if composed is not None:
        t0 = time.perf_counter()
        composed[0]()
        td = time.perf_counter() - t0
        t0 = time.perf_counter()
        composed[1]()
        tg = time.perf_counter() - t0
        sec = a.iterD * td + tg
        return {"sec_per_step": sec, "steps": 1, "value": units_per_step(a, b) / sec, "cores": cores, "batch": b,
                "how": "one D step (%.1f s) and one G step (%.1f s) timed once each, no warm-up; train step = %d x D + G"
                       % (td, tg, a.iterD)}
    for _ in range(n_warm):
        one()
    times = []
    t_all = time.perf_counter()
    for _ in range(max(1, n_timed)):
        t0 = time.perf_counter()
        one()
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_all > budget_s:
            break
    sec = sum(times) / len(times)
    return {"sec_per_step": sec, "steps": len(times), "value": units_per_step(a, b) / sec, "cores": cores, "batch": b,
            "how": "%d step(s) timed, %d warm-up" % (len(times), n_warm)}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = cpu_reference(a, a.steps, min(a.warmup, 1), a.cpu_budget_s)
    metric, unit = metric_of(a)
    sample = "%s at batch %d, fp32, stock torch layers on the host CPU (oracle/ref_*.py); %s" % (
        workload_text(a), r["batch"], r["how"])
    line = {
        "impl": "reference", "metric": metric, "value": r["value"], "unit": unit,
        "n_gpus": a.gpus, "steps": r["steps"], "warmup": min(a.warmup, 1), "ms_per_step": r["sec_per_step"] * 1e3,
        "steps_per_sec": 1.0 / r["sec_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_text(a), "per_gpu_batch": r["batch"], "global_batch": r["batch"],
                   "filterG": a.width, "filterD": a.width, "z_size": a.z, "iterD": a.iterD, "device": "cpu"},
        "cpu_baseline": {"value": r["value"], "unit": unit, "cores": r["cores"], "kind": "port", "sample": sample},
        "e2e": {"value": r["value"], "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------- GPU arm
def conv_flops(key):
    pas, tr, n, xc, xd, yc, yd, k, s, p, dl = key
    small = xd if tr else yd
    return 2.0 * n * small ** 3 * xc * yc * k ** 3


class GanWorkload:
    """dcgan / biggan: b200conv3d.train.GanStep over the product models (the public API)."""

    def __init__(self, a, dev, rank, world, stock=False):
        import torch
        import torch.distributed as dist
        import b200conv3d as B
        self.a, self.dev, self.B = a, dev, B
        torch.manual_seed(1234 + rank)
        if stock:                                   # witness: the same tables from stock torch layers (cuDNN)
            import torch.nn as tnn
            if a.workload == "dcgan":
                netG = B.dcgan.Generator(a.z, a.width, layers=tnn).to(dev)
                netD = B.dcgan.Discriminator(a.width, 128, layers=tnn).to(dev)
            else:
                sys.path.insert(0, os.path.join(ROOT, "oracle"))
                import ref_biggan
                netG, netD = ref_biggan.Generator(a.z, a.width).to(dev), ref_biggan.Discriminator(a.width).to(dev)
        elif a.workload == "dcgan":
            netG, netD = B.dcgan.build(a.z, a.width, a.width, 128, dev)
        else:
            netG, netD = B.biggan.build(a.z, a.width, a.width, dev)
        red_g = red_d = None
        if world > 1 and not stock:                 # identical replicas (DataParallel semantics)
            for p in list(netG.parameters()) + list(netD.parameters()) + list(netG.buffers()) + list(netD.buffers()):
                dist.broadcast(p.data, 0)
            B.functional.invalidate_packs()
            red_g, red_d = B.parallel.GradReducer(netG), B.parallel.GradReducer(netD)
        self.step = B.train.GanStep(netG, netD, a.z, hinge=a.workload == "biggan", iterD=a.iterD, reducer_g=red_g,
                                    reducer_d=red_d, seed=99 + rank,
                                    batch_real_fake=not a.two_pass_d and not stock)
        b = a.batch
        self.npool = 2 * a.iterD                    # distinct real batches, 268 MB each at batch 32: larger than L2 in total
        self.host = None if stock else [(torch.rand(b, 1, 128, 128, 128) * 2 - 1).pin_memory() for _ in range(self.npool)]
        self.resident = [h.to(dev) for h in self.host] if not stock else \
            [torch.rand(b, 1, 128, 128, 128, device=dev) * 2 - 1 for _ in range(self.npool)]
        self.h2d = a.iterD * b * VOX * 4
        self.d2h = (2 * a.iterD + 1) * 4
        self.feed = None

    def step_resident(self, i):
        a = self.a
        self.step.step(self.resident[(i * a.iterD + j) % self.npool] for j in range(a.iterD))

    def step_e2e(self, i):
        # the public API end to end: pinned host batches -> DevicePrefetcher (H2D copy of every batch inside the timed
        # region, overlapped with the running step) -> GanStep.D_step / G_step -> the step's losses read back to the
        # host (trainer.py:251,281 `.item()`)
        import torch
        if self.feed is None:
            def gen():
                i_ = 0
                while True:
                    yield self.host[i_ % self.npool]
                    i_ += 1
            self.feed = self.B.train.DevicePrefetcher(gen(), self.dev)
        losses = []
        for _ in range(self.a.iterD):
            er, ef = self.step.D_step(next(self.feed))
            losses += [er, ef]
        eg, _ = self.step.G_step(self.a.batch)
        losses.append(eg)
        return torch.stack([l.float() for l in losses]).cpu().tolist()

    def finish(self):
        pass

    def nets(self):
        return [self.step.netG, self.step.netD]

    def reducer(self):
        return self.step.reducer_d

    def extra_config(self):
        return {"d_step": ("D(real) and D(fake) of trainer.py:222-241 run as one pass over the concatenated batch "
                           "(no batch-coupled or per-call state in D: same per-sample arithmetic)"
                           if self.step.batch_real_fake else "D(real), D(fake) as two passes (trainer.py:239-241)")}


class StyleGan2Workload:
    """stylegan2: the trainer's StyleGAN2 branches (trainer.py:214-220,262-269) over b200conv3d.stylegan."""

    def __init__(self, a, dev, rank, world, stock=False):
        import torch
        import torch.distributed as dist
        import b200conv3d as B
        self.a, self.dev = a, dev
        p = argparse.Namespace(z_size=a.z, filterG=a.width, filterD=a.width, device=dev, stylegan2=True, stylegan=False)
        torch.manual_seed(1234 + rank)
        self.red_g = self.red_d = None
        if stock:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import ref_stylegan as R
            D, G = R.Discriminator(a.width).to(dev), R.Generator(a.z, a.width).to(dev)
            pl = torch.zeros([], device=dev)
            self.step_D = lambda s, real, z: R.step_D(G, D, s, real, z)
            self.step_G = lambda s, z: R.step_G(G, D, s, z, pl)
        else:
            D, G = B.stylegan.Discriminator(p), B.stylegan.Generator(p)
            loss = B.stylegan.StyleGAN2Loss(p, G.mapping, G.synthesis, D, style_mixing_prob=0.9)
            G, D = G.to(dev), D.to(dev)
            loss.pl_mean = loss.pl_mean.to(dev)
            self.step_D, self.step_G = loss.step_D, loss.step_G
            B.functional.watch_optimizers()
            if world > 1:
                for q in list(G.parameters()) + list(D.parameters()) + list(G.buffers()) + list(D.buffers()):
                    dist.broadcast(q.data, 0)
                B.functional.invalidate_packs()
                self.red_g, self.red_d = B.parallel.GradReducer(G), B.parallel.GradReducer(D)
        self.G, self.D = G, D
        self.optD = torch.optim.Adam(D.parameters(), lr=1e-4, betas=(0.0, 0.9), fused=True)
        self.optG = torch.optim.Adam(G.parameters(), lr=5e-5, betas=(0.0, 0.9), fused=True)
        for q in list(D.parameters()) + list(G.parameters()):
            q.requires_grad = False
        b = a.batch
        self.gen = torch.Generator(device=dev).manual_seed(7 + rank)
        self.npool = 2 * a.iterD
        self.host = None if stock else [(torch.rand(b, 1, 128, 128, 128) * 2 - 1).pin_memory() for _ in range(self.npool)]
        self.resident = [h.to(dev) for h in self.host] if not stock else \
            [torch.rand(b, 1, 128, 128, 128, device=dev) * 2 - 1 for _ in range(self.npool)]
        self.h2d, self.d2h = a.iterD * b * VOX * 4, (2 * a.iterD + 1) * 4
        self.feed = None

    def _noise(self):
        import torch
        return torch.randn(self.a.batch, self.a.z, 1, 1, 1, device=self.dev, generator=self.gen)

    def _d_phase(self, real):
        import torch
        for q in self.D.parameters():
            q.requires_grad = True
        self.D.zero_grad()
        if self.red_d is not None:
            self.red_d.begin(backwards=2)       # step_D calls backward() for the generated and for the real logits
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = self.step_D(1, real, self._noise())
        if self.red_d is not None:
            self.red_d.step(self.optD)
        else:
            self.optD.step()
        for q in self.D.parameters():
            q.requires_grad = False
        return out

    def _g_phase(self):
        import torch
        for q in self.G.parameters():
            q.requires_grad = True
        self.G.zero_grad()
        if self.red_g is not None:
            self.red_g.begin()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = self.step_G(1, self._noise())
        if self.red_g is not None:
            self.red_g.step(self.optG)
        else:
            self.optG.step()
        for q in self.G.parameters():
            q.requires_grad = False
        return out

    def step_resident(self, i):
        for j in range(self.a.iterD):
            self._d_phase(self.resident[(i * self.a.iterD + j) % self.npool])
        self._g_phase()

    def step_e2e(self, i):
        import torch
        import b200conv3d as B
        if self.feed is None:
            def gen():
                i_ = 0
                while True:
                    yield self.host[i_ % self.npool]
                    i_ += 1
            self.feed = B.train.DevicePrefetcher(gen(), self.dev)
        losses = []
        for _ in range(self.a.iterD):
            lr, lg = self._d_phase(next(self.feed))
            losses += [lr, lg]
        lG, _ = self._g_phase()
        losses.append(lG)
        return torch.stack([l.float().reshape(()) for l in losses]).cpu().tolist()

    def finish(self):
        pass

    def nets(self):
        return [self.G, self.D]

    def reducer(self):
        return self.red_d

    def extra_config(self):
        return {"losses": "StyleGAN2Loss.step_D / step_G (stylegan.py:39-99), style_mixing_prob 0.9, step index 1"}


class FidWorkload:
    """fid: eval_utils.fid_3d's feature pass (eval_utils.py:41-44) sharded over the ranks."""

    def __init__(self, a, dev, rank, world, stock=False):
        import torch
        import b200conv3d as B
        self.a, self.dev, self.world, self.B = a, dev, world, B
        torch.manual_seed(0)
        if stock:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import ref_fid_resnet
            self.model = ref_fid_resnet.resnet50().to(dev)
        else:
            self.model = B.fid_resnet.resnet50().to(dev)
        gen = torch.Generator().manual_seed(100 + rank)
        b = a.batch
        self.host = None if stock else [torch.randn(b, 1, 128, 128, 128, generator=gen).pin_memory() for _ in range(2)]
        self.resident = [h.to(dev) for h in self.host] if not stock else [torch.randn(b, 1, 128, 128, 128, device=dev) for _ in range(2)]
        self.h2d, self.d2h = b * VOX * 4, b * 2048 * 4
        self.feats = []
        self.feed = None

    def _features(self, x):
        import torch
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            return self.model(x).float().mean(dim=(2, 3, 4))

    def step_resident(self, i):
        self.feats.append(self._features(self.resident[i % 2]))

    def step_e2e(self, i):
        if self.feed is None:
            def gen():
                i_ = 0
                while True:
                    yield self.host[i_ % 2]
                    i_ += 1
            self.feed = self.B.train.DevicePrefetcher(gen(), self.dev)
        f = self._features(next(self.feed))
        self.feats.append(f)
        return f.cpu()

    def finish(self):
        """eval_metrics.py gathers the per-GPU features of DataParallel on GPU 0: one all_gather of [n, 2048]."""
        import torch
        import torch.distributed as dist
        if not self.feats:
            return
        f = torch.cat(self.feats)
        self.feats = []
        if self.world > 1:
            out = [torch.empty_like(f) for _ in range(self.world)]
            dist.all_gather(out, f)
            f = torch.cat(out)
        self.last_shape = list(f.shape)

    def extra_config(self):
        return {"features": "[steps x batch x gpus, 2048] fp32, all_gather inside the timed region",
                "weights": "random init (MedicalNet resnet_50.pth is not available offline)"}


def make_workload(a, dev, rank, world, stock=False):
    cls = {"dcgan": GanWorkload, "biggan": GanWorkload, "stylegan2": StyleGan2Workload, "fid": FidWorkload}[a.workload]
    return cls(a, dev, rank, world, stock)


# ncu `--set full` captures of the dominant launch of each workload (dram__bytes_read.sum + dram__bytes_write.sum)
TRAFFIC = {
    ("dcgan", 32, 128): (3.272e9, "profiles/r02_prof_conv_DL2fwd_summary.txt (tc_conv_kernel<2, false>, CTA pairs): "
                                  "dram__bytes_read.sum 2.743 GB + dram__bytes_write.sum 0.528 GB per launch "
                                  "(algorithmic 2.69 GB); L2 -> SM 25.8 GB (34.4 GB before the pairs)"),
}


def run_ours(a):
    import torch
    import torch.distributed as dist
    import b200conv3d as B  # noqa: F401
    from b200conv3d import _lib, functional as BF

    BF.STATS[0] = {"all": True, "none": False}.get(a.fused_stats, a.fused_stats)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # NCCL prints its version banner / debug lines to STDOUT by default; the contract is ONE JSON line there
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        sys.stdout.flush()
        saved = os.dup(1)                     # belt and braces: fd 1 -> stderr while the communicator comes up
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    lib = _lib.lib()                                   # raises if libb2c.so is missing
    wl = make_workload(a, dev, rank, world)
    b = a.batch
    metric, unit = metric_of(a)
    units = units_per_step(a, b)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0 = lib.b2c_launch_count()
        e0.record()
        for i in range(k):
            fn(i)
        wl.finish()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item(), lib.b2c_launch_count() - n0

    for i in range(a.warmup):
        wl.step_resident(i)
    wl.finish()
    sampler = ClockSampler(local)
    sampler.start()
    ms, launches = timed(wl.step_resident, a.steps)
    clocks = sampler.result()
    ms_step = ms / a.steps
    value = world * units / (ms_step * 1e-3)

    e2e = None
    if not a.no_e2e:
        wl.step_e2e(0)
        wl.finish()
        ms_e, _ = timed(wl.step_e2e, a.steps)
        e2e = {"value": world * units / (ms_e / a.steps * 1e-3), "unit": unit,
               "ms_per_step": ms_e / a.steps, "h2d_bytes_per_step": wl.h2d, "d2h_bytes_per_step": wl.d2h}

    # per-launch profile of one extra step: every conv launch bracketed by CUDA events
    BF.PROFILE = []
    wl.step_resident(0)
    wl.finish()
    torch.cuda.synchronize()
    prof, BF.PROFILE = BF.PROFILE, None
    agg = {}
    for key, e0, e1, algo in prof:
        t = e0.elapsed_time(e1)
        r = agg.setdefault(key, {"ms": 0.0, "n": 0, "algo": algo})
        r["ms"] += t
        r["n"] += 1
    dense = [(k, r) for k, r in agg.items() if r["algo"] == 2 and min(k[3], k[5]) > 1]
    tc_ms = sum(r["ms"] for k, r in dense)
    tc_fl = sum(conv_flops(k) * r["n"] for k, r in dense)
    tc_n = sum(r["n"] for k, r in dense)
    conv_ms = sum(r["ms"] for r in agg.values())
    peaks = {}
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peaks = json.load(open(pk))
    peak = peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1400.0))
    ach = tc_fl / (tc_ms * 1e-3) / 1e12 if tc_ms > 0 else 0.0
    roofline = {"bound": "tensor", "kernel": "tc_conv_kernel/tc_wgrad_kernel (tcgen05 implicit GEMM)",
                "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak if peak else None,
                "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained" if peaks else "fallback",
                "launches": tc_n, "avg_launch_ms": tc_ms / max(tc_n, 1),
                "flops_per_launch": tc_fl / max(tc_n, 1), "conv_share_of_step": conv_ms / ms_step,
                "traffic": None}
    # the single heaviest launch signature of the step; for dcgan (D.L2 fwd = Conv3d(128->256, k4, s2) at 64^3, with its
    # mirror G.L5 52 % of the step's FLOPs) an `ncu --set full` capture of this very launch is committed
    if dense:
        if a.workload == "dcgan":
            dom = [(k, r) for k, r in dense if k[0] == "fwd" and not k[1] and (k[3], k[5], k[4], k[7], k[8]) ==
                   (a.width, 2 * a.width, 64, 4, 2)]
            dom.sort(key=lambda kr: kr[0][2] != b)     # the per-GPU-batch launch (G step's D pass) is the one captured with ncu
        else:
            dom = sorted(dense, key=lambda kr: -kr[1]["ms"])
        if dom:
            k, r = dom[0]
            fl = conv_flops(k)
            tr = TRAFFIC.get((a.workload, k[2], a.width), (None, None))
            roofline["dominant_launch"] = {
                "signature": "conv%s3d %s %d->%d k%d s%d @%d^3 -> %d^3, batch %d" % ("T" if k[1] else "", k[0], k[3], k[5],
                                                                                      k[7], k[8], k[4], k[6], k[2]),
                "flops_per_launch": fl, "avg_launch_ms": r["ms"] / r["n"], "launches": r["n"],
                "achieved": fl * r["n"] / (r["ms"] * 1e-3) / 1e12, "frac": fl * r["n"] / (r["ms"] * 1e-3) / 1e12 / peak,
                "traffic": tr[0], "traffic_source": tr[1]}
            roofline["traffic"] = tr[0]
    if a.profile_json and rank == 0:
        rows = [{"pass": k[0], "transposed": k[1], "n": k[2], "cin": k[3], "in": k[4], "cout": k[5], "out": k[6],
                 "k": k[7], "stride": k[8], "calls": r["n"], "ms_total": r["ms"], "algo": r["algo"],
                 "tflops": conv_flops(k) * r["n"] / (r["ms"] * 1e-3) / 1e12} for k, r in agg.items()]
        rows.sort(key=lambda r: -r["ms_total"])
        json.dump({"ms_per_step": ms_step, "conv_ms_profiled_step": conv_ms, "rows": rows},
                  open(a.profile_json, "w"), indent=1)
    extra = wl.extra_config()
    if world > 1 and hasattr(wl, "nets"):
        # data-parallel invariants after the timed steps: every replica holds the same, finite parameters (the
        # averaging is deterministic and identical on all ranks), and which transport averaged the gradients
        chk = torch.stack([torch.stack([q.detach().double().sum(), q.detach().double().abs().sum()])
                           for net in wl.nets() for q in net.parameters()]).sum(0)
        allc = [torch.empty_like(chk) for _ in range(world)]
        dist.all_gather(allc, chk)
        extra["replicas_identical"] = bool(all(torch.equal(c, allc[0]) for c in allc)) and \
            bool(torch.isfinite(chk).all().item())
        red = wl.reducer()
        extra["gradient_averaging"] = ("b2c_peer_allreduce_avg over NVLink peer memory, %d blocks" % red.comm.CTAS
                                       if red is not None and red.comm is not None else "NCCL all-reduce (AVG)")
        if _lib.lib().b2c_peer_error() != 0:
            raise SystemExit("bench.py: a peer all-reduce timed out (b2c_peer_error)")
    del wl
    torch.cuda.empty_cache()

    witness = None
    want_wit = a.witness if a.witness is not None else (a.workload == "dcgan" and world == 1)
    if want_wit and rank == 0 and world == 1:
        # "what you get today on this box": the same step on stock torch layers (ATen -> cuDNN) under bf16 autocast
        ww = make_workload(a, dev, rank, 1, stock=True)
        nw, nt = min(a.warmup, 2), min(a.steps, 3)
        for i in range(nw):
            ww.step_resident(i)
        ww.finish()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(nt):
            ww.step_resident(i)
        ww.finish()
        e1.record()
        torch.cuda.synchronize()
        wms = e0.elapsed_time(e1) / nt
        witness = {"ms_per_step": wms, "value": units / (wms * 1e-3), "unit": unit, "steps": nt, "warmup": nw,
                   "impl": "stock torch %s layers (ATen -> cuDNN %s) under bf16 autocast, same GPU, same step schedule"
                           % (torch.__version__, ".".join(str(v) for v in _cudnn_version(torch)))}
        del ww
        torch.cuda.empty_cache()

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        r = cpu_reference(a, 1, 0, a.cpu_budget_s)
        cpu = {"value": r["value"], "unit": unit, "cores": r["cores"], "kind": "port",
               "sample": "%s at batch %d, fp32, stock torch layers on the host CPU (oracle/ref_*.py); %s; %.1f s per step"
                         % (workload_text(a), r["batch"], r["how"], r["sec_per_step"])}
    if rank == 0:
        config = {"workload": workload_text(a), "per_gpu_batch": b, "global_batch": b * world, "filterG": a.width,
                  "filterD": a.width, "z_size": a.z, "iterD": a.iterD, "parallelism": "dp%d" % world,
                  "l2": "inputs larger than L2: distinct input batches of %d MB rotate, per-step activations > 10 GB"
                        % (b * VOX * 4 >> 20)}
        config.update(extra)
        if a.fused_stats != "all":
            config["fused_stats"] = a.fused_stats
        line = {
            "metric": metric, "value": value, "unit": unit, "n_gpus": world,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_step, "steps_per_sec": 1e3 / ms_step,
            "higher_is_better": True, "scaling": "strong" if a.global_batch else "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic", "config": config,
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu,
            "witness": witness,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def _cudnn_version(torch):
    v = torch.backends.cudnn.version() or 0
    return (v // 10000, v // 100 % 100, v % 100)


def main():
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
