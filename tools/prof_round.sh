mkdir -p gpurun_out
set -x
timeout 400 python bench.py --profile-json gpurun_out/r01b_step_profile.json > gpurun_out/r01b_bench.json 2> gpurun_out/r01b_bench.err || exit 1
timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/r01b_bench_s1.json 2>/dev/null || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r01b_launches_bench_steps1.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/r01b_ncu_bench.log 2>&1
python tools/agg_launches.py gpurun_out/r01b_launches_bench_steps1.csv 3 40 > gpurun_out/r01b_launches_summary.txt
timeout 200 python tools/layer_bench.py --config dcgan --batch 32 --iters 1 --only D.L2 --passes fwd > gpurun_out/r01b_layer_DL2.log 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_conv_kernel -s 2 -c 1 -o gpurun_out/r01b_prof_conv_DL2fwd python tools/layer_bench.py --config dcgan --batch 32 --iters 1 --only D.L2 --passes fwd > gpurun_out/r01b_ncu_DL2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fir_sep_kernel -s 20 -c 1 -o gpurun_out/r01b_prof_fir_sep python tools/stylegan_bench.py --steps 1 --warmup 1 > gpurun_out/r01b_ncu_fir.log 2>&1
ls -la gpurun_out/*.ncu-rep
cat gpurun_out/r01b_bench.json
