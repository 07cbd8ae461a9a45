# swapped 128 x 256 tiles (Cout <= 128 on UMMA M) against direct 256 x 128 CTA-pair tiles for the 128-channel outputs of DCGAN
mkdir -p gpurun_out
for cfg in "default:" "noswap:B2C_SWAP_MIN=129"; do
  name=${cfg%%:*}; envs=${cfg#*:}
  echo "=== $name ($envs)"
  env $envs timeout 300 python tools/layer_bench.py --config dcgan --batch 32 --iters 7 --only "G.L5" --passes fwd,dgrad,wgrad 2>&1 | grep "G.L5"
  env $envs timeout 300 python tools/layer_bench.py --config dcgan --batch 32 --iters 7 --only "D.L2" --passes fwd,dgrad 2>&1 | grep "D.L2"
  env $envs timeout 300 python tools/layer_bench.py --config dcgan --batch 64 --iters 5 --only "D.L2" --passes dgrad 2>&1 | grep "D.L2"
done
