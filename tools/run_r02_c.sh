mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_fused_stats.py -q -m gpu -x > gpurun_out/fused_stats.log 2>&1; echo rc=$? >> gpurun_out/fused_stats.log
B2C_FULLSIZE_JSON=gpurun_out/r02_fullsize_parity.json timeout 600 python -m pytest tests/test_gpu_fullsize.py -q -m gpu -s > gpurun_out/fullsize.log 2>&1; echo rc=$? >> gpurun_out/fullsize.log
timeout 200 python tools/step_timeline.py --out gpurun_out/step_timeline.json > gpurun_out/step_timeline.log 2>&1
timeout 400 python bench.py --steps 10 --warmup 3 --profile-json gpurun_out/step_profile.json > gpurun_out/bench_dcgan.json 2> gpurun_out/bench_dcgan.err; echo rc=$? >> gpurun_out/bench_dcgan.err
for f in fused_stats fullsize; do echo "== $f"; tail -n 6 gpurun_out/$f.log; done
grep -n "^E  " gpurun_out/fullsize.log | head
tail -n 30 gpurun_out/step_timeline.log
tail -n 2 gpurun_out/bench_dcgan.err; cut -c1-300 gpurun_out/bench_dcgan.json
