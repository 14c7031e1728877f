#!/bin/bash
mkdir -p gpurun_out
B="python bench.py --workload c5 --steps 5 --warmup 2 --no-cpu-baseline --no-e2e --no-api"
run() { echo "== $1"; env $1 $B 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
t=d['pipeline_roofline']['top_kernels_ms_per_step']
print(round(d['ms_per_step'],2), {k:t[k] for k in ('K_Threshold','K_MinEdge','K_BasinAccum','K_MaskPaint','K_Expand')})"; }
run "X=1"
run "ULK_THR_U=8"
run "ULK_THR_U=2"
run "ULK_U_MinEdge=2"
run "ULK_U_BasinAccum=2"
run "ULK_U_MaskPaint=2"
run "ULK_U_MaskPaint=1"
