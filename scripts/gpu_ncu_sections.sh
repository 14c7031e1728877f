#!/bin/bash
# ncu stall / memory sections for the latency-bound list kernels at a workload far larger than L2
mkdir -p gpurun_out
WL=${1:-c4}
PAT=${2:-'K_BasinAccum|K_K3Union|K_CclUnion|K_MinEdge|K_CompressCand|K_Attach|K_RagInsert'}
CMD="python bench.py --workload $WL --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
timeout 1500 ncu --section SpeedOfLight --section MemoryWorkloadAnalysis --section WarpStateStats --section SchedulerStats --section Occupancy --section LaunchStats \
  --clock-control none --import-source on --kernel-name-base demangled -k regex:"$PAT" -s ${3:-12} -c ${4:-12} -f -o gpurun_out/sect_$WL $CMD > gpurun_out/ncu_sect_$WL.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/ncu_sect_$WL.log | cut -c1-300
