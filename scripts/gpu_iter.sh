#!/bin/bash
# quick iteration on the GPU box: parity tests, then the C5 bench (no CPU arm), optional C3
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --workload c5 --no-cpu-baseline > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "bench c5 rc=$?"
tail -3 gpurun_out/bench_c5.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_c5.json'))
print('ms/step',d['ms_per_step'],'value',d['value'],'e2e',d.get('e2e',{}).get('value'),'kernel ms',d['pipeline_roofline']['kernel_time_ms_per_step'], 'launches', d['gpu_launches'], 'ws', d['workspace_bytes'])
for k,v in d['pipeline_roofline']['top_kernels_ms_per_step'].items(): print(f'{k:50s} {v:8.3f}')
PY
