#!/bin/bash
# ncu --set full of the top voxel kernels at C5 (one bench step after one warm-up), with source;
# the report stays on the box (too large), its raw / source pages come back as CSV
mkdir -p gpurun_out
PAT=${PAT:-'K_BasinAccum|K_MinEdge|K_MaskPaint|ulk_expand_kernel|K_CclWords|K_CompressCand|K_Attach|K_LabelCc|K_FlattenCount|K_PairIoU|ulk_front'}
SRC=${SRC:-'K_BasinAccum K_MinEdge K_MaskPaint ulk_expand_kernel K_CompressCand K_Attach K_CclWords'}
CMD="python bench.py --workload c5 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-api"
timeout 1500 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"$PAT" -c 80 -f \
  -o /tmp/full_c5 $CMD > gpurun_out/ncu_full_c5.log 2>&1; echo "ncu full rc=$?"
tail -2 gpurun_out/ncu_full_c5.log | cut -c1-300
ncu -i /tmp/full_c5.ncu-rep --page raw --csv > gpurun_out/full_c5_raw.csv 2> gpurun_out/raw.err
for k in $SRC; do
  ncu -i /tmp/full_c5.ncu-rep --page source --csv --kernel-name-base demangled -k regex:"$k" > gpurun_out/src_$k.csv 2>> gpurun_out/raw.err
done
ls -la gpurun_out
