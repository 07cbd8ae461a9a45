mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
O="--steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-witness"
CUDA_VISIBLE_DEVICES=0 timeout 300 python bench.py $O 2> gpurun_out/n1.err | cut -c1-250
p=29600
for c in 96 128 148; do
  p=$((p+1))
  B2C_PEER_CTAS=$c timeout 400 $TR --master-port $p bench.py --gpus 2 $O 2> gpurun_out/peer_ctas$c.err | cut -c1-250; echo "ctas=$c rc=$?"
done
timeout 300 $TR --master-port 29521 tools/step_timeline.py --profile-steps 3 --out gpurun_out/step_timeline_2gpu_3steps.json > gpurun_out/step_timeline_2gpu_3steps.log 2>&1; echo rc=$?
