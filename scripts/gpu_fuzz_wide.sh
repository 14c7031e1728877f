#!/bin/bash
# parity fuzz on the GPU: shapes that span several front-end bricks + a fresh seed range of the plain family
mkdir -p gpurun_out
timeout 900 python scripts/gpu_fuzz.py 300 9000 wide > gpurun_out/gpu_fuzz_wide.log 2>&1; echo "wide rc=$?"; tail -3 gpurun_out/gpu_fuzz_wide.log
timeout 900 python scripts/gpu_fuzz.py 400 12000 > gpurun_out/gpu_fuzz_plain.log 2>&1; echo "plain rc=$?"; tail -3 gpurun_out/gpu_fuzz_plain.log
