# N = 8 on one box: N = 1 reference of the same box first, then the peer transport, NCCL, strong scaling, StyleGAN2
mkdir -p gpurun_out
N=${N:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
O="--steps 10 --warmup 3 --no-cpu-baseline --no-witness"
CUDA_VISIBLE_DEVICES=0 timeout 300 python bench.py $O --no-e2e > gpurun_out/bench_dcgan_1gpu_samebox.json 2> gpurun_out/n1.err; echo "n1 rc=$?"; cut -c1-250 gpurun_out/bench_dcgan_1gpu_samebox.json
run() {  # name, port, env, args...
  name=$1; port=$2; envs=$3; shift 3
  env $envs timeout 420 $TR --master-port $port bench.py --gpus $N "$@" > gpurun_out/bench_${name}_${N}gpu.json 2> gpurun_out/bench_${name}_${N}gpu.err
  echo "== $name rc=$?"; grep -v "^\[\|^\*\|OMP_NUM\|^$" gpurun_out/bench_${name}_${N}gpu.err | tail -n 3; cut -c1-250 gpurun_out/bench_${name}_${N}gpu.json; echo
}
run dcgan 29531 "B2C_REDUCE=peer" $O
run dcgan_nccl 29532 "B2C_REDUCE=nccl" $O --no-e2e
run dcgan_strong 29533 "B2C_REDUCE=peer" $O --global-batch 32 --no-e2e
run stylegan2 29534 "B2C_REDUCE=peer" --workload stylegan2 --steps 4 --warmup 3 --no-witness --no-cpu-baseline
