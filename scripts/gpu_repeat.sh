#!/bin/bash
mkdir -p gpurun_out
for i in ${REPS:-1 2 3 4}; do
  timeout 600 python bench.py --workload c5 --no-cpu-baseline --no-e2e --steps ${1:-4} > gpurun_out/rep_$i.json 2> gpurun_out/rep_$i.err
  python - gpurun_out/rep_$i.json <<'PY'
import json,sys
d=json.load(open(sys.argv[1]))
print('ms/step',round(d['ms_per_step'],2),'kernel ms',round(d['pipeline_roofline']['kernel_time_ms_per_step'],2), d['ms_steps_rank0'], d['clocks'])
PY
done
nproc; uptime
