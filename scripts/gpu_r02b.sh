#!/bin/bash
# r02: parity tests + full bench line (all legs)
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 1200 python bench.py --workload c5 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "bench c5 rc=$?"
tail -5 gpurun_out/bench_c5.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_c5.json'))
print('ms/step',d['ms_per_step'],'value',d['value'],'kernel ms',d['pipeline_roofline']['kernel_time_ms_per_step'], 'launches', d['gpu_launches'], 'ws', d['workspace_bytes'])
for k in ('e2e','e2e_u8f16','h2d_ceiling','e2e_api','cpu_baseline'): print(k, json.dumps(d.get(k))[:900])
PY
