#!/bin/bash
# The kernel functors compiled for the host (tests/emu) with AddressSanitizer + UBSan: the emulation
# allocates every scratch buffer with malloc, so an out-of-range index in a functor is an ASan
# report.  Runs the kernel-logic / host-API tests and a short random parity fuzz under it.  CPU only.
set -e
cd "$(dirname "$0")/.."
SRC=$(python -c "from ultrack_b200.csrc.build import SOURCES, HERE; print(' '.join('-x c++ ' + str(HERE / s) for s in SOURCES))")
g++ -O1 -g -std=c++17 -DULK_EMU -fPIC -shared -ffp-contract=off -fsanitize=address,undefined -fno-omit-frame-pointer \
    -o /tmp/libulk_emu_asan.so $SRC
export ULK_EMU_LIB=/tmp/libulk_emu_asan.so ASAN_OPTIONS=detect_leaks=0:halt_on_error=0 UBSAN_OPTIONS=print_stacktrace=1
export LD_PRELOAD=$(gcc -print-file-name=libasan.so):$(gcc -print-file-name=libubsan.so)
python -m pytest tests/test_kernel_logic_emu.py tests/test_host_api.py -q -p no:cacheprovider 2>&1 | tee /tmp/emu_sanitize.log | tail -2
python scripts/emu_fuzz.py ${1:-60} 40000 2>&1 | tee -a /tmp/emu_sanitize.log | tail -1
python scripts/emu_link_fuzz.py ${2:-40} 5000 2>&1 | tee -a /tmp/emu_sanitize.log | tail -1
echo "sanitizer reports: $(grep -c 'runtime error\|ERROR: AddressSanitizer' /tmp/emu_sanitize.log)"
