# what the driver runs at round end on one B200, plus the evidence lines of the round
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/final_gpu_tests.log 2>&1; echo "tests rc=$?"; tail -n 6 gpurun_out/final_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/final_smoke.log
timeout 600 python bench.py --profile-json gpurun_out/r02_step_profile_final.json > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/final_bench.json

