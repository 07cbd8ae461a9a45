mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 python -m pytest tests/test_peer_reduce.py -q -m gpu -x > gpurun_out/peer_test.log 2>&1; echo rc=$? >> gpurun_out/peer_test.log
tail -n 5 gpurun_out/peer_test.log
timeout 400 $TR --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_dcgan_2gpu.json 2> gpurun_out/bench_dcgan_2gpu.err; echo rc=$? >> gpurun_out/bench_dcgan_2gpu.err
grep -v "^\[" gpurun_out/bench_dcgan_2gpu.err | tail -n 6; cut -c1-330 gpurun_out/bench_dcgan_2gpu.json
timeout 300 $TR --master-port 29521 tools/step_timeline.py --out gpurun_out/step_timeline_2gpu.json > gpurun_out/step_timeline_2gpu.log 2>&1; echo rc=$? >> gpurun_out/step_timeline_2gpu.log
grep -v "^\[" gpurun_out/step_timeline_2gpu.log | tail -n 48
B2C_REDUCE=nccl timeout 400 $TR --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --no-e2e > gpurun_out/bench_dcgan_2gpu_nccl.json 2> gpurun_out/bench_dcgan_2gpu_nccl.err; echo rc=$? >> gpurun_out/bench_dcgan_2gpu_nccl.err
grep -v "^\[" gpurun_out/bench_dcgan_2gpu_nccl.err | tail -n 3; cut -c1-330 gpurun_out/bench_dcgan_2gpu_nccl.json
timeout 400 $TR --master-port 29513 bench.py --gpus 2 --workload stylegan2 --steps 3 --warmup 3 --no-witness > gpurun_out/bench_stylegan2_2gpu.json 2> gpurun_out/bench_stylegan2_2gpu.err; echo rc=$? >> gpurun_out/bench_stylegan2_2gpu.err
grep -v "^\[" gpurun_out/bench_stylegan2_2gpu.err | tail -n 3; cut -c1-330 gpurun_out/bench_stylegan2_2gpu.json
