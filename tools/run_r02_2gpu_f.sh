mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29521 tools/step_timeline.py --profile-steps 3 --out gpurun_out/step_timeline_2gpu_3steps.json > gpurun_out/step_timeline_2gpu_3steps.log 2>&1; echo rc=$? >> gpurun_out/step_timeline_2gpu_3steps.log
grep -v "^\[" gpurun_out/step_timeline_2gpu_3steps.log | grep -v "^ *\[\|^ *\]\|^  *[0-9.]*,$\|^ *\"" | head -40
