mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29521 tools/step_timeline.py --out gpurun_out/step_timeline_2gpu.json > gpurun_out/step_timeline_2gpu.log 2>&1; echo rc=$? >> gpurun_out/step_timeline_2gpu.log
timeout 400 $TR --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_dcgan_2gpu.json 2> gpurun_out/bench_dcgan_2gpu.err; echo rc=$? >> gpurun_out/bench_dcgan_2gpu.err
grep -v "^\[" gpurun_out/step_timeline_2gpu.log | tail -n 75
tail -n 3 gpurun_out/bench_dcgan_2gpu.err; cut -c1-330 gpurun_out/bench_dcgan_2gpu.json
