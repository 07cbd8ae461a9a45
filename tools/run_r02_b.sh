mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_fused_stats.py -q -m gpu -x > gpurun_out/fused_stats.log 2>&1; echo rc=$? >> gpurun_out/fused_stats.log
timeout 600 python -m pytest tests -q -m gpu -x --deselect tests/test_gpu_fullsize.py > gpurun_out/gpu_all.log 2>&1; echo rc=$? >> gpurun_out/gpu_all.log
timeout 300 python -m pytest tests/test_gpu_fullsize.py -q -m gpu -k "dcgan_train_step or G.L2 or G.L5 or D.L2" > gpurun_out/fullsize_step.log 2>&1; echo rc=$? >> gpurun_out/fullsize_step.log
timeout 200 python tools/step_timeline.py --out gpurun_out/step_timeline.json > gpurun_out/step_timeline.log 2>&1
timeout 400 python bench.py --steps 10 --warmup 3 --profile-json gpurun_out/step_profile.json > gpurun_out/bench_dcgan.json 2> gpurun_out/bench_dcgan.err; echo rc=$? >> gpurun_out/bench_dcgan.err
for f in fused_stats gpu_all fullsize_step; do echo "== $f"; tail -n 4 gpurun_out/$f.log; done
tail -n 45 gpurun_out/step_timeline.log
tail -n 2 gpurun_out/bench_dcgan.err; cut -c1-700 gpurun_out/bench_dcgan.json
