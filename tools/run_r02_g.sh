mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_stylegan_ops.py -q -m gpu -x -k fir > gpurun_out/fir_tests.log 2>&1; echo rc=$? >> gpurun_out/fir_tests.log; tail -n 4 gpurun_out/fir_tests.log
timeout 300 python tools/stylegan_bench.py --steps 3 --warmup 2 --kernels > gpurun_out/stylegan_bench_g.log 2>&1; grep -n "train_step\|kernels of one call" gpurun_out/stylegan_bench_g.log | head -5; grep -n "fir_\|act_kernel\|d2s" gpurun_out/stylegan_bench_g.log | head -12
timeout 300 python tools/biggan_bench.py --kernels > gpurun_out/biggan_bench_g.log 2>&1; grep -n "kernels of one\|ms   " gpurun_out/biggan_bench_g.log | head -30
