mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/r01k_pytest.log 2>&1; tail -3 gpurun_out/r01k_pytest.log
python tools/layer_bench.py --config dcgan --batch 32 --iters 3 --only D.L5 --passes wgrad 2>&1 | tail -3
python tools/layer_bench.py --config dcgan --batch 32 --iters 3 --only D.L4 --passes wgrad 2>&1 | tail -2
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"tc_wgrad_kernel|wgrad_reduce" -s 2 -c 2 -o gpurun_out/r01k_prof_wgrad_DL5 python tools/layer_bench.py --config dcgan --batch 32 --iters 1 --only D.L5 --passes wgrad > gpurun_out/r01k_ncu.log 2>&1
ls -la gpurun_out/*.ncu-rep
python bench.py --no-cpu-baseline > gpurun_out/r01k_bench.json 2> gpurun_out/r01k_bench.err; cut -c1-330 gpurun_out/r01k_bench.json
