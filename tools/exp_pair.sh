# tile-shape experiments on the heaviest DCGAN layers (isolated launches, L2 flushed between them)
mkdir -p gpurun_out
for cfg in "default:" "pair:B2C_PAIR=1" "pair_mt1:B2C_PAIR=1 B2C_MT1=1" "mt1:B2C_MT1=1"; do
  name=${cfg%%:*}; envs=${cfg#*:}
  echo "=== $name ($envs)"
  env $envs timeout 300 python tools/layer_bench.py --config dcgan --batch 32 --iters 7 --only "L2 conv" --passes fwd,dgrad 2>&1 | grep -v "^layer"
  env $envs timeout 300 python tools/layer_bench.py --config dcgan --batch 32 --iters 7 --only "G.L5" --passes fwd,dgrad 2>&1 | grep "G.L5"
  env $envs timeout 300 python tools/layer_bench.py --config dcgan --batch 32 --iters 7 --only "D.L4" --passes fwd,dgrad 2>&1 | grep "D.L4"
done
