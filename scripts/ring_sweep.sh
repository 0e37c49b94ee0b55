#!/bin/bash
# GPU box: e2e (ExtractorRing) for a few chunk sizes / depths
mkdir -p gpurun_out; rm -f gpurun_out/ring_sweep.log
for v in ${RING:-"96 2" "64 2" "128 2" "192 2" "96 3" "48 3" "128 3"}; do
  set -- $v
  python bench.py --steps 60 --warmup 3 --no-cpu-baseline --no-match --ring-chunk $1 --ring-depth $2 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read())
e=d['e2e']
print('chunk $1 depth $2 value', round(d['value']), 'e2e', round(e['value']), 'ring', round(e['ring_value']), 'single', round(e['single_call_value']), 'ceiling', round(e['h2d_copy_only_frames_per_s']))" | tee -a gpurun_out/ring_sweep.log
done
