#!/bin/bash
# The 8-GPU questions in one call: topology, the H2D/D2H concurrency matrix (1/2/4/8 GPUs copying at once), the torchrun
# bench at N = 8, one process driving all GPUs through BSW_CUDA_DEVICES (C5 stand-in), the multi-device parity test.
cd "$(dirname "$0")/.."
OUT=${1:-gpurun_out/r02_8gpu}; mkdir -p $OUT
nvidia-smi topo -m > $OUT/topo.txt 2>&1
(nproc; lscpu | grep -i "model name\|socket\|numa\|thread\|core(s)"; free -g | head -2) > $OUT/host.txt 2>&1
make -C tools bin/pcie_probe bin/stream_c5 > /dev/null 2>&1
tools/bin/pcie_probe --gpus 8 --pairs 4000000 > $OUT/pcie_matrix.json 2> $OUT/pcie_matrix.err
python -m pytest tests -m gpu -x -q -k "multi_device" > $OUT/pytest_multi.log 2>&1; tail -2 $OUT/pytest_multi.log
for N in 8 2; do
  timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 --reads 1500000 \
    > $OUT/bench_${N}gpu.json 2> $OUT/bench_${N}gpu.err
  tail -c 400 $OUT/bench_${N}gpu.json; tail -2 $OUT/bench_${N}gpu.err
done
python bench.py --gpus 1 --steps 3 --warmup 3 --reads 1500000 > $OUT/bench_1gpu.json 2> $OUT/bench_1gpu.err
for G in 8 1; do
  timeout 120 python bench.py --workload stream_c5 --gpus $G --threads 32,64 --task-reads 5120 --reads 2000000 > $OUT/stream_${G}gpu.json 2> $OUT/stream_${G}gpu.err
  tail -2 $OUT/stream_${G}gpu.err
done
