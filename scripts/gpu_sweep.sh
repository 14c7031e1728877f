#!/bin/bash
# bench runs with env overrides ("A=1,B=2" per argument), prints selected kernel times
mkdir -p gpurun_out
i=0
for kv in "$@"; do
  i=$((i+1))
  env ${kv//,/ } timeout 600 python bench.py --workload c5 --no-cpu-baseline --no-e2e --steps 4 > gpurun_out/sweep_$i.json 2> gpurun_out/sweep_$i.err
  python - gpurun_out/sweep_$i.json "$kv" <<'PY'
import json,sys
d=json.load(open(sys.argv[1]))
t=d['pipeline_roofline']['top_kernels_ms_per_step']
print(sys.argv[2], 'ms/step',round(d['ms_per_step'],2), ' '.join(f"{k[2:]}={t.get(k)}" for k in ['K_BasinAccum','K_MaskPaint','K_CclUnion','K_MinEdge','K_Hook1','K_RagInsert','K_K3Union','K_Threshold']))
PY
done
