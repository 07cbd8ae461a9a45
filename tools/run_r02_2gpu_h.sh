mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
O="--steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-witness"
for g in 0 1; do CUDA_VISIBLE_DEVICES=$g timeout 300 python bench.py $O 2> gpurun_out/n1_$g.err | cut -c100-200; done
timeout 400 $TR --master-port 29601 bench.py --gpus 2 $O > gpurun_out/bench_dcgan_2gpu_peer.json 2> gpurun_out/bench_dcgan_2gpu_peer.err; echo "peer rc=$?"
python -c "
import json;d=json.loads(open('gpurun_out/bench_dcgan_2gpu_peer.json').read());print(d['ms_per_step'], d['config'].get('replicas_identical'), d['config'].get('gradient_averaging'), d['gpu_launches'])"
grep -v "^\[\|^\*\|OMP_NUM\|^$" gpurun_out/bench_dcgan_2gpu_peer.err | tail -n 5
timeout 400 $TR --master-port 29602 bench.py --gpus 2 --workload stylegan2 --steps 3 --warmup 3 --no-witness --no-cpu-baseline --no-e2e > gpurun_out/bench_stylegan2_2gpu.json 2> gpurun_out/bench_stylegan2_2gpu.err; echo "sg2 rc=$?"
python -c "
import json;d=json.loads(open('gpurun_out/bench_stylegan2_2gpu.json').read());print(d['ms_per_step'], d['config'].get('replicas_identical'), d['config'].get('gradient_averaging'))"
grep -v "^\[\|^\*\|OMP_NUM\|^$" gpurun_out/bench_stylegan2_2gpu.err | tail -n 5
timeout 300 $TR --master-port 29521 tools/step_timeline.py --profile-steps 3 --out gpurun_out/step_timeline_2gpu_3steps.json > gpurun_out/step_timeline_2gpu_3steps.log 2>&1; echo rc=$?
