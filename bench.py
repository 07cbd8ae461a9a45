#!/usr/bin/env python
"""bench.py — the driver's measurement contract for the 3D-GAN conv3d hot path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[1]): 3D DCGAN, 128^3 single-channel volumes, batch 32
per GPU, bf16, one *train step* = iterD(=2) x D_step + 1 x G_step exactly as
trainer.py:296-303 schedules them; data-parallel over N GPUs with
NCCL gradient all-reduce overlapped with backward.  Synthetic data, random-init weights.

Prints ONE JSON line (rank 0).  `value` = generated voxels per second with the real
batches already resident in HBM; `e2e` = the same step fed from pinned HOST memory with
the loss scalars read back (.item()) as the reference trainer does.
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


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=32, help="per-GPU batch")
    ap.add_argument("--iterD", type=int, default=2)
    ap.add_argument("--width", type=int, default=128, help="filterG = filterD (reference default 128)")
    ap.add_argument("--z", type=int, default=512)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cpu-batch", type=int, default=2)
    ap.add_argument("--cpu-budget-s", type=float, default=150.0)
    ap.add_argument("--profile-json", default="", help="write the per-launch conv timing table here")
    return ap.parse_args()


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


# ------------------------------------------------------------------ CPU reference arm
def cpu_reference_steps(a, n_timed, n_warm, budget_s):
    """The reference's own CPU path (stock torch layers; oracle/ref_dcgan.py restates
    dcgan.py + trainer.py and is pinned against the imported reference by
    tests/golden) on the host cores: batch `cpu_batch`, fp32, one train step =
    iterD x d_step + g_step.  Runs `n_warm` untimed and up to `n_timed` timed steps, but
    stops early once `budget_s` seconds of timed work have elapsed (>= 1 step)."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ref_dcgan
    cores = physical_cores()
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    g = ref_dcgan.Generator(a.z, a.width)
    d = ref_dcgan.Discriminator(a.width)
    tr = ref_dcgan.RefTrainer(g, d, a.z)
    b = a.cpu_batch

    def one():
        for _ in range(a.iterD):
            real = torch.rand(b, 1, 128, 128, 128) * 2 - 1
            tr.d_step(real)
        tr.g_step(b)

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
    return {"sec_per_step": sec, "steps": len(times), "value": b * VOX / sec, "cores": cores, "batch": b,
            "torch_threads": torch.get_num_threads()}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = cpu_reference_steps(a, a.steps, min(a.warmup, 1), a.cpu_budget_s)
    sample = ("%d train step(s) (iterD=%d D_step + 1 G_step) of DCGAN 128^3 width %d at batch %d fp32 on the "
              "host CPU, %d warm-up; stock torch layers (oracle/ref_dcgan.py)" %
              (r["steps"], a.iterD, a.width, r["batch"], min(a.warmup, 1)))
    line = {
        "impl": "reference", "metric": "dcgan128_train_voxels_per_sec", "value": r["value"], "unit": "voxels/s",
        "n_gpus": a.gpus, "steps": r["steps"], "warmup": min(a.warmup, 1), "ms_per_step": r["sec_per_step"] * 1e3,
        "steps_per_sec": 1.0 / r["sec_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "3D DCGAN 128^3 1-channel, train step = %d x D_step + G_step (trainer.py:296-303)"
                               % a.iterD, "per_gpu_batch": r["batch"], "global_batch": r["batch"],
                   "filterG": a.width, "filterD": a.width, "z_size": a.z, "iterD": a.iterD, "device": "cpu"},
        "cpu_baseline": {"value": r["value"], "unit": "voxels/s", "cores": r["cores"], "kind": "port",
                         "sample": sample},
        "e2e": {"value": r["value"], "unit": "voxels/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------- GPU arm
def conv_flops(key):
    pas, tr, n, xc, xd, yc, yd, k, s, p, dl = key
    small = xd if tr else yd
    return 2.0 * n * small ** 3 * xc * yc * k ** 3


def run_ours(a):
    import torch
    import torch.distributed as dist
    import b200conv3d as B
    from b200conv3d import _lib, functional as BF

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
    torch.manual_seed(1234 + rank)
    netG, netD = B.dcgan.build(a.z, a.width, a.width, 128, dev)
    if world > 1:                                      # identical replicas (DataParallel semantics)
        for p in list(netG.parameters()) + list(netD.parameters()) + list(netG.buffers()) + list(netD.buffers()):
            dist.broadcast(p.data, 0)
    red_g = B.parallel.GradReducer(netG) if world > 1 else None
    red_d = B.parallel.GradReducer(netD) if world > 1 else None
    step = B.train.GanStep(netG, netD, a.z, iterD=a.iterD, reducer_g=red_g, reducer_d=red_d, seed=99 + rank)

    b = a.batch
    npool = 2 * a.iterD                                # distinct real batches, 268 MB each (fp32) > L2 in total
    host = [(torch.rand(b, 1, 128, 128, 128) * 2 - 1).pin_memory() for _ in range(npool)]
    resident = [h.to(dev) for h in host]
    h2d_per_step = a.iterD * host[0].numel() * 4
    d2h_per_step = (2 * a.iterD + 1) * 4

    def step_resident(i):
        step.step(resident[(i * a.iterD + j) % npool] for j in range(a.iterD))

    def host_batches():
        i = 0
        while True:
            yield host[i % npool]
            i += 1

    feed = [None]

    def step_e2e(i):
        # the public API end to end: pinned host batches -> DevicePrefetcher (H2D copy of every
        # batch inside the timed region, overlapped with the running step) -> GanStep.D_step /
        # G_step -> the step's losses read back to the host (trainer.py:251,281 `.item()`)
        if feed[0] is None:
            feed[0] = B.train.DevicePrefetcher(host_batches(), dev)
        losses = []
        for j in range(a.iterD):
            er, ef = step.D_step(next(feed[0]))
            losses += [er, ef]
        eg, _ = step.G_step(b)
        losses.append(eg)
        return torch.stack([l.float() for l in losses]).cpu().tolist()

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
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item(), lib.b2c_launch_count() - n0

    for i in range(a.warmup):
        step_resident(i)
    sampler = ClockSampler(local)
    sampler.start()
    ms, launches = timed(step_resident, a.steps)
    clocks = sampler.result()
    ms_step = ms / a.steps
    value = world * b * VOX / (ms_step * 1e-3)

    e2e = None
    if not a.no_e2e:
        step_e2e(0)
        ms_e, _ = timed(step_e2e, a.steps)
        e2e = {"value": world * b * VOX / (ms_e / a.steps * 1e-3), "unit": "voxels/s",
               "ms_per_step": ms_e / a.steps, "h2d_bytes_per_step": h2d_per_step, "d2h_bytes_per_step": d2h_per_step}

    # per-launch profile of one extra step: every conv launch bracketed by CUDA events
    BF.PROFILE = []
    step_resident(0)
    torch.cuda.synchronize()
    prof, BF.PROFILE = BF.PROFILE, None
    agg = {}
    for key, e0, e1, algo in prof:
        t = e0.elapsed_time(e1)
        r = agg.setdefault(key, {"ms": 0.0, "n": 0, "algo": algo})
        r["ms"] += t
        r["n"] += 1
    tc_ms = sum(r["ms"] for k, r in agg.items() if r["algo"] == 2 and min(k[3], k[5]) > 1)
    tc_fl = sum(conv_flops(k) * r["n"] for k, r in agg.items() if r["algo"] == 2 and min(k[3], k[5]) > 1)
    tc_n = sum(r["n"] for k, r in agg.items() if r["algo"] == 2 and min(k[3], k[5]) > 1)
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
    # the single heaviest launch signature (D.L2 fwd = Conv3d(128->256, k4, s2) at 64^3: with its mirror G.L5 52 % of
    # the step's FLOPs), for which an `ncu --set full` capture of this very launch is committed
    dom = [(k, r) for k, r in agg.items() if k[0] == "fwd" and not k[1] and (k[3], k[5], k[4], k[7], k[8]) ==
           (a.width, 2 * a.width, 64, 4, 2)]
    dom.sort(key=lambda kr: kr[0][2] != b)     # the per-GPU-batch launch (G step's D pass) is the one captured with ncu;
    if dom:                                    # the D steps run real + fake as one batch-2b pass
        k, r = dom[0]
        fl = conv_flops(k)
        roofline["dominant_launch"] = {
            "signature": "conv3d fwd %d->%d k4 s2 @64^3, batch %d (dcgan.py:165)" % (k[3], k[5], k[2]),
            "flops_per_launch": fl, "avg_launch_ms": r["ms"] / r["n"], "launches": r["n"],
            "achieved": fl * r["n"] / (r["ms"] * 1e-3) / 1e12, "frac": fl * r["n"] / (r["ms"] * 1e-3) / 1e12 / peak,
            "traffic": 2.842e9 if (k[2], a.width) == (32, 128) else None,
            "traffic_source": "profiles/r01_prof_conv_DL2fwd_summary.txt: dram__bytes_read.sum 2.316 GB + "
                              "dram__bytes_write.sum 0.525 GB (algorithmic 2.69 GB)"}
    if a.profile_json and rank == 0:
        rows = [{"pass": k[0], "transposed": k[1], "n": k[2], "cin": k[3], "in": k[4], "cout": k[5], "out": k[6],
                 "k": k[7], "stride": k[8], "calls": r["n"], "ms_total": r["ms"], "algo": r["algo"],
                 "tflops": conv_flops(k) * r["n"] / (r["ms"] * 1e-3) / 1e12} for k, r in agg.items()]
        rows.sort(key=lambda r: -r["ms_total"])
        json.dump({"ms_per_step": ms_step, "conv_ms_profiled_step": conv_ms, "rows": rows},
                  open(a.profile_json, "w"), indent=1)

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        r = cpu_reference_steps(a, 1, 0, a.cpu_budget_s)
        cpu = {"value": r["value"], "unit": "voxels/s", "cores": r["cores"], "kind": "port",
               "sample": "1 train step (iterD=%d) of the same DCGAN 128^3 width-%d model at batch %d, fp32, stock "
                         "torch layers on the host CPU (oracle/ref_dcgan.py), no warm-up, %.1f s"
                         % (a.iterD, a.width, r["batch"], r["sec_per_step"])}
    if rank == 0:
        line = {
            "metric": "dcgan128_train_voxels_per_sec", "value": value, "unit": "voxels/s", "n_gpus": world,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_step, "steps_per_sec": 1e3 / ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "3D DCGAN 128^3 1-channel, train step = %d x D_step + G_step "
                                   "(trainer.py:296-303), BASELINE configs[1]" % a.iterD,
                       "per_gpu_batch": b, "global_batch": b * world, "filterG": a.width, "filterD": a.width,
                       "z_size": a.z, "iterD": a.iterD, "parallelism": "dp%d" % world,
                       "l2": "inputs larger than L2: %d distinct real batches of %d MB rotate, per-step "
                             "activations > 20 GB" % (npool, host[0].numel() * 4 >> 20),
                       "d_step": ("D(real) and D(fake) of trainer.py:222-241 run as one pass over the concatenated "
                                  "batch (LayerNorm D: same per-sample arithmetic)" if step.batch_real_fake
                                  else "D(real), D(fake) as two passes")},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
