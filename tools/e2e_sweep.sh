#!/bin/bash
# sweep of the host pipeline knobs (upload-ahead depth, chunk size): prints e2e GCUPS and ms/step
for a in 1 2; do for c in 262144 524288 1048576; do
  BSW_CUDA_AHEAD=$a BSW_CUDA_CHUNK_PAIRS=$c python bench.py --steps 3 --warmup 2 --cpu-sample 20000 2>&1 | tail -1 > /tmp/l.json
  python -c "import json; d=json.load(open('/tmp/l.json')); print('ahead $a chunk $c e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],1))"
done; done
