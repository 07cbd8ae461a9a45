# N = 4 check of the data-parallel path (the W = 4 instantiation of the averaging kernel with real processes)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
O="--steps 8 --warmup 3 --no-cpu-baseline --no-witness --no-e2e"
CUDA_VISIBLE_DEVICES=0 timeout 300 python bench.py $O > gpurun_out/bench_dcgan_1gpu_samebox4.json 2> gpurun_out/n1.err; echo "n1 rc=$?"
timeout 400 $TR --master-port 29801 bench.py --gpus 4 $O > gpurun_out/bench_dcgan_4gpu.json 2> gpurun_out/bench_dcgan_4gpu.err; echo "peer rc=$?"
B2C_REDUCE=nccl timeout 400 $TR --master-port 29802 bench.py --gpus 4 $O > gpurun_out/bench_dcgan_nccl_4gpu.json 2> gpurun_out/bench_dcgan_nccl_4gpu.err; echo "nccl rc=$?"
for f in bench_dcgan_1gpu_samebox4 bench_dcgan_4gpu bench_dcgan_nccl_4gpu; do python - <<P
import json
d=json.loads(open("gpurun_out/$f.json").read())
print("$f", round(d["ms_per_step"],2), d["config"].get("replicas_identical"), d["config"].get("gradient_averaging"), d["clocks"])
P
done
grep -v "^\[\|^\*\|OMP_NUM\|^$\|NCCL version" gpurun_out/bench_dcgan_4gpu.err | tail -n 3
