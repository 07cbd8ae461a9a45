mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_stylegan_ops.py tests/test_stylegan_net.py -q -m gpu -x > gpurun_out/sg_tests.log 2>&1; echo rc=$? >> gpurun_out/sg_tests.log; tail -n 4 gpurun_out/sg_tests.log
timeout 600 python -m pytest tests/test_gpu_golden.py tests/test_gpu_fullsize.py -q -m gpu -x -k "biggan or BigGAN" > gpurun_out/biggan_tests.log 2>&1; echo rc=$? >> gpurun_out/biggan_tests.log; tail -n 4 gpurun_out/biggan_tests.log
timeout 600 python bench.py --workload stylegan2 --steps 3 --warmup 3 --no-cpu-baseline --no-witness > gpurun_out/bench_stylegan2_f.json 2> gpurun_out/bench_stylegan2_f.err; echo rc=$?; cut -c90-220 gpurun_out/bench_stylegan2_f.json
timeout 600 python bench.py --workload biggan --steps 3 --warmup 3 --no-cpu-baseline --no-witness > gpurun_out/bench_biggan_f.json 2> gpurun_out/bench_biggan_f.err; echo rc=$?; cut -c90-220 gpurun_out/bench_biggan_f.json
timeout 300 python tools/stylegan_bench.py --steps 2 --warmup 2 > gpurun_out/stylegan_bench_f.log 2>&1; grep -n "fir\|kernels of one call" gpurun_out/stylegan_bench_f.log | head
