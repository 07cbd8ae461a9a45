mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_window_conv.py -q -m gpu -x > gpurun_out/window.log 2>&1; echo rc=$? >> gpurun_out/window.log
timeout 300 python tools/narrow_bench.py --json gpurun_out/narrow_bench.json > gpurun_out/narrow_bench.log 2>&1; echo rc=$? >> gpurun_out/narrow_bench.log
timeout 120 python -m pytest tests/test_gpu_fullsize.py -q -m gpu -k "statistics_at_headline" > gpurun_out/fullsize_stats.log 2>&1; echo rc=$? >> gpurun_out/fullsize_stats.log
timeout 600 python bench.py --workload biggan --steps 3 --warmup 2 --witness --profile-json gpurun_out/biggan_profile.json > gpurun_out/bench_biggan.json 2> gpurun_out/bench_biggan.err; echo rc=$? >> gpurun_out/bench_biggan.err
timeout 600 python bench.py --workload stylegan2 --steps 3 --warmup 2 --no-cpu-baseline --profile-json gpurun_out/stylegan2_profile.json > gpurun_out/bench_stylegan2.json 2> gpurun_out/bench_stylegan2.err; echo rc=$? >> gpurun_out/bench_stylegan2.err
for f in window fullsize_stats; do echo "== $f"; tail -n 4 gpurun_out/$f.log; done
cat gpurun_out/narrow_bench.log
tail -n 3 gpurun_out/bench_biggan.err; cut -c1-260 gpurun_out/bench_biggan.json
tail -n 3 gpurun_out/bench_stylegan2.err; cut -c1-260 gpurun_out/bench_stylegan2.json
