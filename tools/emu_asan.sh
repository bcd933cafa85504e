#!/bin/bash
# AddressSanitizer over the kernels' per-thread code: builds tests/emu (the __host__ __device__ cores compiled for the host)
# with -fsanitize=address and runs the CPU kernel-logic suite on it.  compute-sanitizer is closed on the GPU pool; this
# covers the row arrays, the register-array indexing and the scratch layouts of every core.  ~40 minutes (the emulated warps
# switch contexts at every shuffle).
set -e
cd "$(dirname "$0")/.."
ASAN=$(g++ -print-file-name=libasan.so)
python -c "import sys; sys.path.insert(0, 'tests'); import emu; emu.lib()"          # make sure the plain build exists
cp tests/emu/libbswemu.so /tmp/libbswemu_plain.so
g++ -O1 -g -std=c++17 -fPIC -shared -fsanitize=address -fno-omit-frame-pointer -o tests/emu/libbswemu.so tests/emu/emu.cpp
trap 'cp /tmp/libbswemu_plain.so tests/emu/libbswemu.so; touch tests/emu/libbswemu.so' EXIT
LD_PRELOAD=$ASAN ASAN_OPTIONS=detect_leaks=0:detect_stack_use_after_return=0 python -m pytest tests/test_kernel_logic_cpu.py -x -q
