#!/bin/bash
mkdir -p gpurun_out
python bench.py --workload c5 --steps 6 --warmup 2 --no-cpu-baseline --no-e2e --no-api 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
t=d['pipeline_roofline']['top_kernels_ms_per_step']
print(round(d['ms_per_step'],2), round(d['pipeline_roofline']['kernel_time_ms_per_step'],2), {k:t[k] for k in ('K_FrontTile','K_Threshold')})"
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:ulk_front_tile -c 3 python bench.py --workload c5 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-api 2>&1 | grep -E "dram__bytes|gpu__time" | tail -3
