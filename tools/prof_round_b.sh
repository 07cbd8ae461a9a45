# second ncu pass of round 1: the StyleGAN glue kernels at BASELINE config 4 (run after the plain command exited 0)
mkdir -p gpurun_out
set -x
timeout 300 python tools/stylegan_bench.py --steps 1 --warmup 1 > gpurun_out/r01c_stylegan_plain.log 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fir_sep_kernel -s 4 -c 1 -o gpurun_out/r01c_prof_fir_sep python tools/stylegan_bench.py --steps 1 --warmup 1 > gpurun_out/r01c_ncu_fir.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mod_epilogue_fwd -s 10 -c 1 -o gpurun_out/r01c_prof_mod_epilogue python tools/stylegan_bench.py --steps 1 --warmup 1 > gpurun_out/r01c_ncu_mod.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_conv_kernel -s 15 -c 1 -o gpurun_out/r01c_prof_conv_fold32 python tools/stylegan_bench.py --steps 1 --warmup 1 > gpurun_out/r01c_ncu_conv.log 2>&1
ls -la gpurun_out/r01c_*.ncu-rep
