for call in 1000 2452 5000 10000 20000; do
 for cfg in "0 0" "4096 8192" "100000 100000" "0 100000" "2000 6000" "8192 20000"; do
  set -- $cfg
  BSW_CUDA_KSWV_L32=$1 BSW_CUDA_KSWV_L16=$2 PROBE_THREADS=1 PROBE_PAIRS=120000 PROBE_CALL=$call python tools/kswv_stream_probe.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print($call, '$1/$2', round(d['us_per_call_per_thread']), d['engine_us_per_call']['flight_us_per_batch'], d['equal'])"
 done
done
