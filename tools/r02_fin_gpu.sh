#!/bin/bash
# round-2 closing GPU call: the allocation-free entry point's test alone, then the whole GPU suite (it runs last there)
mkdir -p gpurun_out/r02_fin
python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/r02_fin/zz_alone.log 2>&1; echo "alone rc=$?"; tail -3 gpurun_out/r02_fin/zz_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/r02_fin/gpu_suite.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/r02_fin/gpu_suite.log
