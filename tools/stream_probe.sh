#!/bin/bash
# C5 stand-in on the GPU box: writes the workloads with the CPU restatement (python), runs tools/bin/stream_c5 over
# thread counts / arena kinds, prints one JSON line per run.  Usage: tools/stream_probe.sh <n_reads> <outdir> [reads per task]
set -e
cd "$(dirname "$0")/.."
N=${1:-400000}; OUT=${2:-gpurun_out/stream}; TR=${3:-512}; mkdir -p $OUT
python - <<PY
import sys; sys.path[:0] = ["oracle", "tests"]
import numpy as np, ora
n = $N
d = ora.gen_bsw(n, seed=0xC5, nthreads=16)
L, lr, lq = d["left"]; R, rr, rq = d["right"]
ora.write_stream_workload("$OUT/bsw.bin", 0, L, R, lr, lq, rr, rq, ora.task_bounds(L["seqid"], n, $TR), ora.task_bounds(R["seqid"], n, $TR))
K, kr, kq = ora.gen_kswv(max(2500, n // 2), seed=0xC7)
tb = np.arange(0, len(K) + 2499, 2500).clip(max=len(K)).astype(np.int64)
ora.write_stream_workload("$OUT/kswv.bin", 1, K, K[:0], kr, kq, kr[:0], kq[:0], tb, np.zeros(len(tb), np.int64))
PY
for T in 1 4 8 16 32; do
  for P in 0 1; do
    tools/bin/stream_c5 --in $OUT/bsw.bin --threads $T --pinned $P --reps 2
  done
done
tools/bin/stream_c5 --in $OUT/bsw.bin --threads 16 --pinned 0 --per-task-handles 0 --reps 2
for T in 1 8 32; do tools/bin/stream_c5 --in $OUT/kswv.bin --threads $T --reps 2; done
rm -f $OUT/bsw.bin $OUT/kswv.bin
