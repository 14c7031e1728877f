#!/bin/bash
# one ncu --set full capture of selected kernels (B200_PROFILING.md recipe); args: pattern, count, workload
mkdir -p gpurun_out
PAT=${1:-'K_MinEdge|K_BasinAccum|K_K3Union|K_CclUnion|K_Hook1|K_CompressCand|K_RagInsert'}
WL=${3:-c3}
CMD="python bench.py --workload $WL --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"$PAT" -s ${4:-20} -c ${2:-14} -f -o gpurun_out/full_$WL $CMD > gpurun_out/ncu_full_$WL.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/ncu_full_$WL.log | cut -c1-300
