# final N = 8 set of the round: all-reduce microbench, every workload (peer transport), saved under gpurun_out/
mkdir -p gpurun_out
N=${N:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29701 tools/peer_bench.py --json gpurun_out/peer_bench_${N}gpu.json 2> gpurun_out/peer_bench_${N}gpu.err | tee gpurun_out/peer_bench_${N}gpu.log
O="--no-cpu-baseline --no-witness"
CUDA_VISIBLE_DEVICES=0 timeout 300 python bench.py $O --no-e2e --steps 10 --warmup 3 > gpurun_out/bench_dcgan_1gpu_samebox.json 2> gpurun_out/n1.err; echo "n1 rc=$?"; cut -c100-200 gpurun_out/bench_dcgan_1gpu_samebox.json
run() {  # name, port, args...
  name=$1; port=$2; shift 2
  timeout 420 $TR --master-port $port bench.py --gpus $N $O "$@" > gpurun_out/bench_${name}_${N}gpu.json 2> gpurun_out/bench_${name}_${N}gpu.err
  echo "== $name rc=$?"; grep -v "^\[\|^\*\|OMP_NUM\|^$\|NCCL version" gpurun_out/bench_${name}_${N}gpu.err | tail -n 3
  python - <<P
import json
d=json.loads(open("gpurun_out/bench_${name}_${N}gpu.json").read())
print(round(d["ms_per_step"],2), d["value"], d["unit"], d["scaling"], d["config"].get("replicas_identical"), d["config"].get("gradient_averaging"), d["clocks"])
P
}
run dcgan 29531 --steps 10 --warmup 3
run dcgan_strong 29533 --steps 10 --warmup 3 --global-batch 32 --no-e2e
run biggan 29534 --workload biggan --steps 4 --warmup 3
run stylegan2 29535 --workload stylegan2 --steps 4 --warmup 3
run fid 29536 --workload fid --steps 50 --warmup 3
B2C_REDUCE=nccl run dcgan_strong_nccl 29537 --steps 10 --warmup 3 --global-batch 32 --no-e2e
