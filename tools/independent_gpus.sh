# 8 INDEPENDENT single-GPU runs at the same time (no communication at all): what each GPU of the box does under a full-box load
mkdir -p gpurun_out
O="--steps 12 --warmup 4 --no-cpu-baseline --no-witness --no-e2e"
for g in 0 1 2 3 4 5 6 7; do
  CUDA_VISIBLE_DEVICES=$g timeout 300 python bench.py $O > gpurun_out/indep_dev$g.json 2> gpurun_out/indep_dev$g.err &
done
wait
for g in 0 1 2 3 4 5 6 7; do python - <<P
import json
d=json.loads(open("gpurun_out/indep_dev$g.json").read())
print("dev$g", round(d["ms_per_step"],2), d["clocks"])
P
done
