mkdir -p gpurun_out
N=${N:-2}   # tools/peer_bench.py under torchrun on N GPUs
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29701 tools/peer_bench.py --json gpurun_out/peer_bench_${N}gpu.json 2> gpurun_out/peer_bench_${N}gpu.err | tee gpurun_out/peer_bench_${N}gpu.log
grep -v "^\[\|^\*\|OMP_NUM\|^$" gpurun_out/peer_bench_${N}gpu.err | tail -n 5
