# ncu --set full captures of the two bandwidth kernels added in round 2 (after their commands ran clean without ncu)
mkdir -p gpurun_out
timeout 200 python tools/layer_bench.py --config dcgan --batch 32 --iters 1 --only G.L6 --passes fwd > gpurun_out/r02_layer_GL6.log 2>&1; echo "layer rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:col2im1_k4s2_tma_kernel -s 1 -c 1 -o gpurun_out/r02_prof_col2im_k4s2 python tools/layer_bench.py --config dcgan --batch 32 --iters 1 --only G.L6 --passes fwd > gpurun_out/r02_ncu_col2im.log 2>&1; echo "ncu col2im rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:fir_tile_kernel -s 1 -c 1 -o gpurun_out/r02_prof_fir_tile python tools/stylegan_bench.py --steps 1 --warmup 1 > gpurun_out/r02_ncu_fir_tile.log 2>&1; echo "ncu fir rc=$?"
ls -la gpurun_out/*.ncu-rep
