#!/usr/bin/env python
"""Kernel timeline of ONE warm DCGAN train step (BASELINE config 2) taken with torch.profiler (CUPTI activity
records, no replay): per-kernel in-step time, GPU idle gaps, host enqueue time.  ncu launch lists time every
kernel alone with a cold cache; this is what the kernels cost inside the power-capped step.

  python tools/step_timeline.py [--out profiles/r02_step_timeline.json]
  torchrun --nproc-per-node 2 ... tools/step_timeline.py   (data-parallel step; rank 0 reports)
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/step_timeline.json")
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--width", type=int, default=128)
    ap.add_argument("--two-pass", action="store_true", help="D(real), D(fake) as two passes (the reference's schedule)")
    ap.add_argument("--profile-steps", type=int, default=1,
                    help="profile this many back-to-back steps (> 1: steady state, no synchronisation in front)")
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    from torch.profiler import ProfilerActivity, profile
    import b200conv3d as B
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:            # under torchrun: the data-parallel step (rank 0 reports)
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(1234)
    g, d = B.dcgan.build(512, a.width, a.width, 128, dev)
    kw = {}
    if world > 1:
        for p_ in list(g.parameters()) + list(d.parameters()):
            dist.broadcast(p_.data, 0)
        B.functional.invalidate_packs()
        kw.update(reducer_g=B.parallel.GradReducer(g), reducer_d=B.parallel.GradReducer(d))
    step = B.train.GanStep(g, d, 512, iterD=2, seed=99 + rank, batch_real_fake=not a.two_pass, **kw)
    reals = [(torch.rand(a.batch, 1, 128, 128, 128, device=dev) * 2 - 1) for _ in range(4)]

    def one(i):
        step.step(reals[(2 * i + j) % 4] for j in range(2))

    for i in range(4):
        one(i)
    torch.cuda.synchronize()
    # host enqueue time vs device time of 3 steps
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for i in range(3):
        one(i)
    e1.record()
    host_ms = (time.perf_counter() - t0) * 1e3 / 3
    torch.cuda.synchronize()
    dev_ms = e0.elapsed_time(e1) / 3
    if a.profile_steps > 1:
        one(3)                                    # the profiled steps start with the GPU busy
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for i in range(a.profile_steps):
            one(i)
        torch.cuda.synchronize()
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    ks = sorted(((e.time_range.start, e.time_range.end, e.name) for e in evs), key=lambda t: t[0])
    t_first, t_last = ks[0][0], max(k[1] for k in ks)
    busy, cur_end, gaps = 0.0, t_first, []
    for s, e, n in ks:
        if s > cur_end:
            gaps.append((s - cur_end, n))
            busy += e - s
            cur_end = e
        elif e > cur_end:
            busy += e - cur_end
            cur_end = e
    # compute kernels only (NCCL runs on its side stream): idle time of the compute stream = exposed communication
    def is_comm(n):
        return "nccl" in n.lower() or "peer_allreduce" in n

    comp = [k for k in ks if not is_comm(k[2])]
    nccl_ms = sum(k[1] - k[0] for k in ks if is_comm(k[2]) and not k[2].startswith("nccl:")) / 1e3
    comm_rows = sorted(((k[1] - k[0]) / 1e3 for k in ks if is_comm(k[2]) and not k[2].startswith("nccl:")), reverse=True)
    cbusy, cend = 0.0, comp[0][0]
    cgaps = []
    for s, e, n in comp:
        if s > cend:
            cgaps.append((s - cend, n))
            cbusy += e - s
            cend = e
        elif e > cend:
            cbusy += e - cend
            cend = e
    cgaps.sort(key=lambda gg: -gg[0])
    by = {}
    for s, e, n in ks:
        n = n.split("(")[0].replace("void ", "")[:70]
        r = by.setdefault(n, [0.0, 0])
        r[0] += (e - s) / 1e3
        r[1] += 1
    rows = sorted(([n, v[0], v[1]] for n, v in by.items()), key=lambda r: -r[1])
    gaps.sort(key=lambda gg: -gg[0])
    out = {"ms_per_step_events": dev_ms, "host_enqueue_ms_per_step": host_ms,
           "profiled_step_span_ms": (t_last - t_first) / 1e3, "gpu_busy_ms": busy / 1e3,
           "gpu_idle_ms": (t_last - t_first - busy) / 1e3, "kernels": len(ks),
           "world": world, "profiled_steps": a.profile_steps, "comm_kernel_ms": nccl_ms,
           "longest_comm_kernels_ms": [round(v, 3) for v in comm_rows[:12]], "compute_busy_ms": cbusy / 1e3,
           "compute_idle_ms": (comp[-1][1] - comp[0][0] - cbusy) / 1e3 if comp else None,
           "largest_compute_gaps_us": [[round(gg[0], 1), gg[1][:60]] for gg in cgaps[:12]],
           "idle_gaps_over_20us": sum(1 for gg in gaps if gg[0] > 20), "largest_gaps_us": [[round(gg[0], 1), gg[1][:60]] for gg in gaps[:15]],
           "by_kernel_ms": rows}
    if rank == 0:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        json.dump(out, open(a.out, "w"), indent=1)
        print(json.dumps({k: v for k, v in out.items() if k != "by_kernel_ms"}, indent=1))
        for n, ms, c in rows[:40]:
            print("%9.3f ms %5d  %s" % (ms, c, n))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
