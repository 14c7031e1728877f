#!/bin/bash
# Round evidence run on one B200: parity tests, smoke, the default bench line (with cpu_baseline
# and e2e), the reference arm, the ncu launch list of the same bench command, per-kernel DRAM
# traffic at C5 (2 metrics, few replays) and one `--set full` capture of the top kernels at C3.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "bench c5 rc=$?"
tail -3 gpurun_out/bench_c5.err
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
CMD="python bench.py --workload c5 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 3000 --csv \
  --log-file gpurun_out/launches_c5.csv $CMD > gpurun_out/ncu_bench_c5.log 2>&1; echo "ncu launches rc=$?"
PAT='K_BasinAccum|K_Threshold|K_K3Union|K_MinEdge|K_CompressCand|K_CclUnion|K_Hook1|K_RagInsert|K_Attach|K_MaskPaint|K_AssignBasin|K_PairIoU'
timeout 1200 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"$PAT" -s 14 -c 14 -f \
  -o gpurun_out/full_c3 python bench.py --workload c3 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full_c3.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out
