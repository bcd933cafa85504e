#!/bin/bash
# compute-sanitizer over one small invocation of every kernel family (__graft_entry__.smoke()).  Usage: tools/sanitize.sh memcheck|racecheck <outdir>
cd "$(dirname "$0")/.."
TOOL=${1:-memcheck}; OUT=${2:-gpurun_out/sanitize}; mkdir -p $OUT
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke_plain.log 2>&1 || { tail -5 $OUT/smoke_plain.log; exit 1; }
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 20 --log-file $OUT/${TOOL}.log python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke_${TOOL}.log 2>&1
echo "exit $?" >> $OUT/smoke_${TOOL}.log
tail -3 $OUT/smoke_${TOOL}.log; grep -c "Invalid\|Race\|hazard" $OUT/${TOOL}.log; tail -5 $OUT/${TOOL}.log
