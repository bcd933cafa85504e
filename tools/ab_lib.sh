#!/bin/bash
# A/B of two builds of libbswcuda.so on one box: tools/ab_lib.sh <experimental .so> <out dir> [rounds]
# Runs the t16 and stream_c5 bench lines alternately with the in-tree library ("base") and the experimental one ("exp").
set -u
EXP=$1; OUT=$2; ROUNDS=${3:-2}
LIB=bwa-mem2-arm_b200/libbswcuda.so
mkdir -p "$OUT"; cp "$LIB" "$OUT/base.so.keep"
for r in $(seq 1 "$ROUNDS"); do
  for v in base exp; do
    if [ $v = exp ]; then cp "$EXP" "$LIB"; else cp "$OUT/base.so.keep" "$LIB"; fi
    python bench.py --workload t16 --steps 3 --warmup 3 2>/dev/null | grep '^{' | tail -1 > "$OUT/t16_${v}_$r.json"
    python bench.py --workload stream_c5 --task-reads 5120 --steps 3 --warmup 3 2>/dev/null | grep '^{' | tail -1 > "$OUT/stream_${v}_$r.json"
    python - "$OUT" $v $r <<'PY'
import json, sys
out, v, r = sys.argv[1:]
for w in ("t16", "stream"):
    try:
        d = json.load(open(f"{out}/{w}_{v}_{r}.json"))
        print(v, r, w, d["value"], d.get("roofline", {}).get("frac"), d.get("parity"))
    except Exception as e:
        print(v, r, w, "ERR", e)
PY
  done
done
cp "$OUT/base.so.keep" "$LIB"; rm -f "$OUT/base.so.keep"
