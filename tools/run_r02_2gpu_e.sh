# per-GPU N = 1 step time next to the N = 2 step time of the same box (peer transport and NCCL)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
O="--steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-witness"
for g in 0 1; do
  CUDA_VISIBLE_DEVICES=$g timeout 300 python bench.py $O > gpurun_out/bench_dcgan_1gpu_dev$g.json 2> gpurun_out/bench_dcgan_1gpu_dev$g.err
  echo "dev$g rc=$?"; cut -c1-250 gpurun_out/bench_dcgan_1gpu_dev$g.json
done
timeout 400 $TR --master-port 29511 bench.py --gpus 2 $O > gpurun_out/bench_dcgan_2gpu_peer.json 2> gpurun_out/bench_dcgan_2gpu_peer.err; echo "peer rc=$?"
cut -c1-250 gpurun_out/bench_dcgan_2gpu_peer.json
B2C_REDUCE=nccl timeout 400 $TR --master-port 29512 bench.py --gpus 2 $O > gpurun_out/bench_dcgan_2gpu_nccl.json 2> gpurun_out/bench_dcgan_2gpu_nccl.err; echo "nccl rc=$?"
cut -c1-250 gpurun_out/bench_dcgan_2gpu_nccl.json
timeout 400 $TR --master-port 29513 bench.py --gpus 2 $O > gpurun_out/bench_dcgan_2gpu_peer2.json 2> gpurun_out/bench_dcgan_2gpu_peer2.err; echo "peer rc=$?"
cut -c1-250 gpurun_out/bench_dcgan_2gpu_peer2.json
