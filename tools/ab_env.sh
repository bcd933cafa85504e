#!/bin/bash
# A/B of an environment knob on one box: tools/ab_env.sh <out dir> <rounds> <workload> VAR=a VAR=b
# Runs `bench.py --workload W` alternately with each setting; prints value, roofline.frac, ms_per_step, e2e.
OUT=$1; ROUNDS=$2; W=$3; shift 3
mkdir -p "$OUT"
for r in $(seq 1 "$ROUNDS"); do
  for kv in "$@"; do
    env "$kv" python bench.py --workload "$W" --steps 5 --warmup 3 --cpu-sample 20000 2>/dev/null | grep '^{' | tail -1 > "$OUT/${W}_${kv}_$r.json"
    python - "$OUT/${W}_${kv}_$r.json" "$kv" $r <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print(sys.argv[2], sys.argv[3], d["value"], d.get("roofline", {}).get("frac"), d["ms_per_step"], d.get("e2e", {}).get("value"), d.get("parity", {}).get("mismatches"))
except Exception as e:
    print(sys.argv[2], sys.argv[3], "ERR", e)
PY
  done
done
