#!/usr/bin/env python
"""b2c_peer_allreduce_avg alone (no compute beside it): time per call and NVLink bytes per second per GPU for a few
bucket sizes and block counts, next to NCCL's all-reduce of the same buffer.  Run under torchrun:

  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/peer_bench.py [--json out.json]

Per rank a call READS (N-1)/N of the bucket from its peers and WRITES (N-1)/N of it to them, and its peers do the same
to it: 2 (N-1)/N bucket bytes cross the rank's NVLink ports in EACH direction (the volume of a ring all-reduce).
`GB/s` below is that per-direction figure (NVLink 5: 900 GB/s per direction per GPU before protocol overhead)."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--json", default="")
    ap.add_argument("--iters", type=int, default=8)
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    dist.init_process_group("nccl", device_id=dev)
    from b200conv3d import parallel
    comm = parallel.PeerComm.get(None, dev)
    if comm is None:
        raise SystemExit("peer transport unavailable")
    rows = []
    for mb in (32, 128, 512):
        n = mb * (1 << 20) // 4
        t, table = comm.bucket(n)
        ref = torch.empty(n, device=dev)

        def timed(fn, stream):
            for _ in range(2):
                fn()
            stream.synchronize()
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(stream):
                e0.record()
                for _ in range(a.iters):
                    fn()
                e1.record()
            stream.synchronize()
            ms = torch.tensor([e0.elapsed_time(e1) / a.iters], device=dev)
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            return ms.item()

        for ctas in (32, 64, 128, 160):
            parallel.PeerComm.CTAS = ctas
            ms = timed(lambda: comm.average(table, n), comm.stream)
            rows.append({"bucket_mb": mb, "impl": "b2c_peer_allreduce_avg", "ctas": ctas, "ms": ms,
                         "gbps_per_direction": 2 * (world - 1) / world * mb / 1024 / (ms * 1e-3)})
        ms = timed(lambda: dist.all_reduce(ref, op=dist.ReduceOp.AVG), torch.cuda.current_stream())
        rows.append({"bucket_mb": mb, "impl": "nccl all_reduce AVG", "ctas": None, "ms": ms,
                     "gbps_per_direction": 2 * (world - 1) / world * mb / 1024 / (ms * 1e-3),
                     "note": "ring-equivalent bytes; NVLS moves about half of them"})
        del t, ref
    if rank == 0:
        for r in rows:
            print("%4d MB  %-24s ctas %-4s %8.3f ms  %7.1f GB/s per direction" %
                  (r["bucket_mb"], r["impl"], r["ctas"], r["ms"], r["gbps_per_direction"]))
        if a.json:
            json.dump({"world": world, "rows": rows}, open(a.json, "w"), indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
