#!/bin/bash
# A/B builds of libbswcuda.so: tools/ab.sh name1 "-DFLAG1=..." name2 "-DFLAG2" ...  ->  bwa-mem2-arm_b200/ab/<name>.so
# (they travel to the GPU box; run them there with tools/ab_bench.py)
set -e
cd "$(dirname "$0")/../bwa-mem2-arm_b200"
mkdir -p ab
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  /usr/local/cuda/bin/nvcc $flags -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -shared -o ab/$name.so csrc/bsw_api.cu -lcudart &
done
wait
ls -la ab
