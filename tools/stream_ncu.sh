#!/bin/bash
# per-launch durations of the staged path: one host thread, C5 workload.  Usage: tools/stream_ncu.sh <n_reads> <outdir> <reads per task>
set -e
cd "$(dirname "$0")/.."
N=${1:-100000}; OUT=${2:-gpurun_out/stream}; TR=${3:-5120}; mkdir -p $OUT
python - <<PY
import sys; sys.path[:0] = ["oracle", "tests"]
import numpy as np, ora
n = $N
d = ora.gen_bsw(n, seed=0xC5, nthreads=16)
L, lr, lq = d["left"]; R, rr, rq = d["right"]
ora.write_stream_workload("$OUT/bsw.bin", 0, L, R, lr, lq, rr, rq, ora.task_bounds(L["seqid"], n, $TR), ora.task_bounds(R["seqid"], n, $TR))
PY
tools/bin/stream_c5 --in $OUT/bsw.bin --threads 1 --reps 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/stream_launches_$TR.csv tools/bin/stream_c5 --in $OUT/bsw.bin --threads 1 --reps 1 > $OUT/ncu_stream.log 2>&1
rm -f $OUT/bsw.bin
