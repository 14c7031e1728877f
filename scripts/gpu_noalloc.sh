#!/bin/bash
# the allocation-free entry point: its test file alone in a fresh process, then the whole GPU suite
# (the file is collected last)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x --timeout 300 > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "alone rc=$?"
tail -15 gpurun_out/pytest_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/pytest_gpu.log 2>&1; echo "suite rc=$?"
tail -5 gpurun_out/pytest_gpu.log
