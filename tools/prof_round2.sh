# round-2 evidence on one B200: fused-statistics A/B, ncu launch list of the bench step, ncu --set full of the dominant launch
mkdir -p gpurun_out
O="--steps 6 --warmup 3 --no-cpu-baseline --no-witness --no-e2e"
for m in all sample none; do
  timeout 300 python bench.py $O --fused-stats $m --profile-json gpurun_out/r02_step_profile_stats_$m.json > gpurun_out/r02_bench_stats_$m.json 2> gpurun_out/r02_bench_stats_$m.err
  echo "stats=$m rc=$?"; cut -c100-200 gpurun_out/r02_bench_stats_$m.json
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches_bench_steps1.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-witness --no-e2e > gpurun_out/r02_ncu_bench.log 2>&1; echo "ncu list rc=$?"
python tools/agg_launches.py gpurun_out/r02_launches_bench_steps1.csv 3 40 > gpurun_out/r02_launches_summary.txt; tail -n 3 gpurun_out/r02_launches_summary.txt
timeout 200 python tools/layer_bench.py --config dcgan --batch 32 --iters 1 --only D.L2 --passes fwd > gpurun_out/r02_layer_DL2.log 2>&1; echo "layer rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_conv_kernel -s 2 -c 1 -o gpurun_out/r02_prof_conv_DL2fwd python tools/layer_bench.py --config dcgan --batch 32 --iters 1 --only D.L2 --passes fwd > gpurun_out/r02_ncu_DL2.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out/*.ncu-rep
