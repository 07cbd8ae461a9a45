mkdir -p gpurun_out
set -x
timeout 400 python bench.py --profile-json gpurun_out/r01p_step_profile.json > gpurun_out/r01p_bench.json 2> gpurun_out/r01p_bench.err || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r01p_launches_bench_steps1.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/r01p_ncu_bench.log 2>&1
python tools/agg_launches.py gpurun_out/r01p_launches_bench_steps1.csv 3 40 > gpurun_out/r01p_launches_summary.txt
head -30 gpurun_out/r01p_launches_summary.txt
cat gpurun_out/r01p_bench.json
