"""BASELINE config 5: FID_ResNet 3-D feature extraction (eval_metrics.py path,
eval_utils.fid_3d, eval_utils.py:41-52) over synthetic 128^3 volumes, sharded over the
ranks of a torchrun launch (one all_gather of the [n,2048] features at the end).

  python tools/fid_bench.py [--volumes 200] [--batch 25]
  python -m torch.distributed.run --nproc-per-node 8 tools/fid_bench.py --volumes 10000

Forward only, bf16 autocast, BatchNorm in train mode (the reference never calls .eval(),
SURVEY F10), random-init weights (resnet_50.pth is not available offline).  Prints one JSON
line with volumes/s and conv TFLOP/s (0.4158 TFLOP per volume).
"""
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
    ap.add_argument("--volumes", type=int, default=200, help="total volumes over all ranks")
    ap.add_argument("--batch", type=int, default=25)
    ap.add_argument("--warmup", type=int, default=2)
    a = ap.parse_args()
    import b200conv3d as B
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # keep stdout for the result line
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(0)
    model = B.fid_resnet.resnet50().to(dev)
    per_rank = a.volumes // world
    nb = max(per_rank // a.batch, 1)
    gen = torch.Generator(device=dev)
    gen.manual_seed(100 + rank)
    pool = [torch.randn(a.batch, 1, 128, 128, 128, device=dev, generator=gen) for _ in range(2)]
    feats = []
    with torch.autocast("cuda", dtype=torch.bfloat16):
        for i in range(a.warmup):
            B.fid_resnet.features(model, pool[i % 2])
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(nb):
            feats.append(B.fid_resnet.features(model, pool[i % 2]))
        f = torch.cat(feats)
        if world > 1:
            out = [torch.empty_like(f) for _ in range(world)]
            dist.all_gather(out, f)
            f = torch.cat(out)
        e1.record()
        torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    vols = nb * a.batch * world
    if rank == 0:
        s = ms.item() * 1e-3
        print(json.dumps({"metric": "fid_resnet50_3d_volumes_per_sec", "value": vols / s, "unit": "volumes/s",
                          "n_gpus": world, "volumes": vols, "batch_per_gpu": a.batch, "seconds": s,
                          "conv_tflops": vols * 0.4158 / s, "dtype": "bf16", "features_shape": list(f.shape),
                          "finite": bool(torch.isfinite(f).all())}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
