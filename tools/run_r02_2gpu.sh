mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_dcgan_2gpu.json 2> gpurun_out/bench_dcgan_2gpu.err; echo rc=$? >> gpurun_out/bench_dcgan_2gpu.err
NCCL_MAX_CTAS=8 timeout 400 $TR --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_dcgan_2gpu_cta8.json 2> gpurun_out/bench_dcgan_2gpu_cta8.err; echo rc=$? >> gpurun_out/bench_dcgan_2gpu_cta8.err
timeout 400 $TR --master-port 29513 bench.py --gpus 2 --steps 10 --warmup 3 --global-batch 32 --no-e2e > gpurun_out/bench_dcgan_2gpu_strong.json 2> gpurun_out/bench_dcgan_2gpu_strong.err; echo rc=$? >> gpurun_out/bench_dcgan_2gpu_strong.err
for f in bench_dcgan_2gpu bench_dcgan_2gpu_cta8 bench_dcgan_2gpu_strong; do echo "== $f"; tail -n 3 gpurun_out/$f.err; cut -c1-330 gpurun_out/$f.json; echo; done
