#!/bin/bash
# parity tests, C5 bench, then extra bench runs with env overrides given as arguments ("VAR=val")
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
show() {
python - "$1" <<'PY'
import json,sys
d=json.load(open(sys.argv[1]))
print('ms/step',d['ms_per_step'],'value',d['value'],'e2e',d.get('e2e',{}).get('value'),'kernel ms',d['pipeline_roofline']['kernel_time_ms_per_step'], 'launches', d['gpu_launches'], 'ws', d['workspace_bytes'])
for k,v in list(d['pipeline_roofline']['top_kernels_ms_per_step'].items())[:int(sys.argv[2]) if len(sys.argv)>2 else 30]: print(f'{k:50s} {v:8.3f}')
PY
}
timeout 600 python bench.py --workload c5 --no-cpu-baseline > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "bench c5 rc=$?"
tail -3 gpurun_out/bench_c5.err
show gpurun_out/bench_c5.json
i=0
for kv in "$@"; do
  i=$((i+1))
  env $kv timeout 600 python bench.py --workload c5 --no-cpu-baseline --no-e2e > gpurun_out/bench_c5_v$i.json 2> gpurun_out/bench_c5_v$i.err; echo "bench $kv rc=$?"
  show gpurun_out/bench_c5_v$i.json
done
