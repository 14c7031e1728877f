#!/bin/bash
# GPU-box check: parity tests, smoke, bench; with "ncu" as first argument also the ncu launch
# list of the same bench command (B200_PROFILING.md recipe).  Every step has its own timeout.
mkdir -p gpurun_out
WL=${ULK_CHECK_WORKLOAD:-c3}
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --workload $WL > gpurun_out/bench_$WL.json 2> gpurun_out/bench_$WL.err; echo "bench $WL rc=$?" | tee -a gpurun_out/bench_$WL.err
tail -c 3000 gpurun_out/bench_$WL.json; tail -5 gpurun_out/bench_$WL.err
if [ "$1" == "ncu" ]; then
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_$WL.csv \
  python bench.py --workload $WL --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_bench_$WL.log 2>&1; echo "ncu rc=$?"
fi
