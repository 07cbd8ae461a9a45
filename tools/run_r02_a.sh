mkdir -p gpurun_out
bash tools/exp_pair.sh > gpurun_out/exp_pair.log 2>&1
timeout 300 python -m pytest tests/test_gpu_fullsize.py -q -m gpu -k dcgan_train_step > gpurun_out/fullsize_step.log 2>&1; echo rc=$? >> gpurun_out/fullsize_step.log
timeout 400 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_dcgan.json 2> gpurun_out/bench_dcgan.err; echo rc=$? >> gpurun_out/bench_dcgan.err
timeout 300 python bench.py --workload fid --steps 8 --warmup 2 --witness > gpurun_out/bench_fid.json 2> gpurun_out/bench_fid.err; echo rc=$? >> gpurun_out/bench_fid.err
timeout 500 python bench.py --workload biggan --steps 3 --warmup 2 --witness > gpurun_out/bench_biggan.json 2> gpurun_out/bench_biggan.err; echo rc=$? >> gpurun_out/bench_biggan.err
timeout 500 python bench.py --workload stylegan2 --steps 3 --warmup 2 --witness > gpurun_out/bench_stylegan2.json 2> gpurun_out/bench_stylegan2.err; echo rc=$? >> gpurun_out/bench_stylegan2.err
tail -3 gpurun_out/*.err gpurun_out/fullsize_step.log
cat gpurun_out/exp_pair.log
