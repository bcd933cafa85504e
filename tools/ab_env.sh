#!/bin/bash
# A/B of environment knobs on one box: tools/ab_env.sh <out dir> <rounds> <workload> "VAR=a" "VAR=b VAR2=c" ...
# Runs `bench.py --workload W` alternately with each setting; prints value, roofline.frac, ms_per_step, e2e, e2e_variants.
OUT=$1; ROUNDS=$2; W=$3; shift 3
mkdir -p "$OUT"
for r in $(seq 1 "$ROUNDS"); do
  for kv in "$@"; do
    tag=$(echo "$kv" | tr ' =' '__')
    env $kv python bench.py --workload "$W" --steps 5 --warmup 3 --cpu-sample 20000 2>/dev/null | grep '^{' | tail -1 > "$OUT/${W}_${tag}_$r.json"
    python - "$OUT/${W}_${tag}_$r.json" "$kv" $r <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    var = {k: round(v["value"], 1) if isinstance(v, dict) and "value" in v else v for k, v in (d.get("e2e_variants") or {}).items()}
    print(sys.argv[2], sys.argv[3], round(d["value"], 1), d.get("roofline", {}).get("frac"), round(d["ms_per_step"], 3),
          round(d.get("e2e", {}).get("value", 0), 1), var, d.get("parity", {}).get("mismatches"))
except Exception as e:
    print(sys.argv[2], sys.argv[3], "ERR", e)
PY
  done
done
