#!/bin/bash
# one ncu --set full capture of selected kernels on the C3 frame (B200_PROFILING.md recipe)
mkdir -p gpurun_out
PAT=${1:-'K_MinEdge|K_BasinAccum|K_K3Union|K_CclUnion|K_Hook1|K_JoinKey|K_Compress|K_RagInsert'}
CMD="python bench.py --workload c3 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"$PAT" -s 20 -c ${2:-14} -f -o gpurun_out/full_c3 $CMD > gpurun_out/ncu_full_c3.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/ncu_full_c3.log
