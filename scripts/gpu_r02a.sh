#!/bin/bash
# r02 baseline evidence at HEAD: parity tests, C5 bench, ncu launch list with DRAM bytes
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --workload c5 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "bench c5 rc=$?"
tail -3 gpurun_out/bench_c5.err
CMD="python bench.py --workload c5 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 3000 --csv \
  --log-file gpurun_out/launches_c5.csv $CMD > gpurun_out/ncu_bench_c5.log 2>&1; echo "ncu launches rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_c5.json'))
print('ms/step',d['ms_per_step'],'value',d['value'],'e2e',d.get('e2e',{}).get('value'),'kernel ms',d['pipeline_roofline']['kernel_time_ms_per_step'], 'launches', d['gpu_launches'], 'ws', d['workspace_bytes'])
for k,v in d['pipeline_roofline']['top_kernels_ms_per_step'].items(): print(f'{k:50s} {v:8.3f}')
PY
