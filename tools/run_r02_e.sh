mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_conv.py -q -m gpu -x > gpurun_out/gpu_conv.log 2>&1; echo rc=$? >> gpurun_out/gpu_conv.log
timeout 600 python bench.py --workload biggan --steps 3 --warmup 2 --no-cpu-baseline --no-witness --profile-json gpurun_out/biggan_profile_e.json > gpurun_out/bench_biggan_e.json 2> gpurun_out/bench_biggan_e.err; echo rc=$? >> gpurun_out/bench_biggan_e.err
tail -n 5 gpurun_out/gpu_conv.log
tail -n 3 gpurun_out/bench_biggan_e.err; cut -c1-300 gpurun_out/bench_biggan_e.json
