# 8-GPU lines of every bench workload (one box, one rank per GPU); each saved under gpurun_out/
mkdir -p gpurun_out
N=${N:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() {  # name, port, args...
  name=$1; port=$2; shift 2
  timeout 420 $TR --master-port $port bench.py --gpus $N "$@" > gpurun_out/bench_${name}_${N}gpu.json 2> gpurun_out/bench_${name}_${N}gpu.err
  echo rc=$? >> gpurun_out/bench_${name}_${N}gpu.err
  echo "== $name"; grep -v "^\[" gpurun_out/bench_${name}_${N}gpu.err | tail -n 3; cut -c1-420 gpurun_out/bench_${name}_${N}gpu.json; echo
}
run dcgan 29531 --steps 10 --warmup 3
run dcgan_strong 29532 --steps 10 --warmup 3 --global-batch 32 --no-e2e
run biggan 29533 --workload biggan --steps 4 --warmup 3 --no-witness
run stylegan2 29534 --workload stylegan2 --steps 4 --warmup 3 --no-witness
run fid 29535 --workload fid --steps 50 --warmup 3 --no-witness
nvidia-smi topo -m > gpurun_out/topo_${N}gpu.txt 2>&1
