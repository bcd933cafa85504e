#!/bin/bash
# Engine settings x host threads on one C5 workload.  Usage: tools/stream_matrix.sh <n_reads> <outdir> <reads per task>
set -e
cd "$(dirname "$0")/.."
N=${1:-2000000}; OUT=${2:-gpurun_out/stream}; TR=${3:-5120}; mkdir -p $OUT
python - <<PY
import sys; sys.path[:0] = ["oracle", "tests"]
import numpy as np, ora
n = $N
d = ora.gen_bsw(n, seed=0xC5, nthreads=16)
L, lr, lq = d["left"]; R, rr, rq = d["right"]
ora.write_stream_workload("$OUT/bsw.bin", 0, L, R, lr, lq, rr, rq, ora.task_bounds(L["seqid"], n, $TR), ora.task_bounds(R["seqid"], n, $TR))
PY
# inflight slots wide_max wide_lat wide_split spin
for CFG in "3 6 32768 8192 0 400" "3 6 32768 32768 0 400" "3 6 32768 2048 0 400" "3 6 65536 8192 0 400" "3 6 32768 8192 5 400" \
           "5 8 32768 8192 0 400" "3 6 0 0 0 400"; do
  set -- $CFG
  for T in 1 8 16 32; do
    echo -n "{\"inflight\": $1, \"slots\": $2, \"wide_max\": $3, \"wide_lat\": $4, \"wide_split\": $5, \"spin\": $6, \"run\": "
    BSW_CUDA_INFLIGHT=$1 BSW_CUDA_STAGE_SLOTS=$2 BSW_CUDA_WIDE_MAX=$3 BSW_CUDA_WIDE_LAT=$4 BSW_CUDA_WIDE_SPLIT=$5 BSW_CUDA_SPIN=$6 \
      tools/bin/stream_c5 --in $OUT/bsw.bin --threads $T --pinned 0 --reps 2 | tr -d '\n'
    echo "}"
  done
done
rm -f $OUT/bsw.bin
