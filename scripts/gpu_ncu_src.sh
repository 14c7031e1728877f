#!/bin/bash
# source-level ncu capture (full set + source counters) of selected kernels at a workload
mkdir -p gpurun_out
WL=${1:-c4}
PAT=${2:-K_BasinAccum}
CMD="python bench.py --workload $WL --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
timeout 1500 ncu --set full --section SourceCounters --clock-control none --import-source on --kernel-name-base demangled -k regex:"$PAT" -s ${3:-2} -c ${4:-2} -f -o gpurun_out/src_$WL $CMD > gpurun_out/ncu_src_$WL.log 2>&1; echo "ncu rc=$?"
tail -2 gpurun_out/ncu_src_$WL.log | cut -c1-200
