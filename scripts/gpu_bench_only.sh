#!/bin/bash
# short C5 bench only (no tests): per-kernel times for an A/B of one kernel
mkdir -p gpurun_out
python bench.py --workload c5 --steps 8 --warmup 3 --no-cpu-baseline --no-e2e --no-api > gpurun_out/bench_quick.json 2>gpurun_out/bench_quick.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_quick.json'))
print('ms/step',round(d['ms_per_step'],3),'kernel ms',round(d['pipeline_roofline']['kernel_time_ms_per_step'],3),'launches',d['gpu_launches'])
t=d['pipeline_roofline']['top_kernels_ms_per_step']
print({k:v for k,v in list(t.items())[:14]})
print(d["ms_steps_rank0"])
PY
