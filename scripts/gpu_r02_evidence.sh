#!/bin/bash
# Round-2 evidence run on one B200: smoke, parity tests, the default bench line (all legs), the
# reference arm, the ncu launch list of the same bench command with DRAM bytes, and one
# `--set full` capture of the top kernels at C5 (raw page exported as CSV on the box).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 1800 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 1200 python bench.py > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "bench c5 rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
CMD="python bench.py --workload c5 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-api"
timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 3000 --csv \
  --log-file gpurun_out/launches_c5.csv $CMD > gpurun_out/ncu_bench_c5.log 2>&1; echo "ncu launches rc=$?"
PAT='K_BasinAccum|K_MinEdge|K_MaskPaint|ulk_front_tile|ulk_threshold|K_CclWords|K_CompressCand|K_Attach|K_LabelCc|K_FlattenCount|K_PairIoU|K_Hook1|K_AssignBasin|K_WriteLabels'
timeout 1500 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"$PAT" -c 90 -f \
  -o /tmp/full_c5 $CMD > gpurun_out/ncu_full_c5.log 2>&1; echo "ncu full rc=$?"
ncu -i /tmp/full_c5.ncu-rep --page raw --csv > gpurun_out/full_c5_raw.csv 2> gpurun_out/raw.err
cuobjdump -sass ultrack_b200/csrc/libultrack_b200.so 2>/dev/null | grep -E "UTMALDG|SYNCS|LDS|STS" | awk '{print $2}' | sort | uniq -c | sort -rn | head -12 > gpurun_out/sass_tma_smem.txt
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_c5.json'))
print('ms/step',d['ms_per_step'],'value',d['value'],'kernel ms',d['pipeline_roofline']['kernel_time_ms_per_step'], 'frac', d['pipeline_roofline']['frac'])
for k in ('e2e','e2e_u8f16','e2e_api','cpu_baseline'): print(k, json.dumps(d.get(k))[:400])
print(open('gpurun_out/bench_ref.json').read()[:600])
PY
ls -la gpurun_out
