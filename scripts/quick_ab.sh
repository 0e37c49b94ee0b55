#!/bin/bash
# GPU box: extraction parity file, then the stage clocks of the default build (twice).
mkdir -p gpurun_out
python -m pytest tests/test_gpu_extract.py tests/test_gpu_prep.py -x -q 2>&1 | tail -2 | tee gpurun_out/quick_ab.log
for i in 1 2; do
  python bench.py --steps 50 --warmup 3 --no-cpu-baseline --no-match 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('value', round(d['value']), 'e2e', round(d['e2e']['value']), d['roofline']['stage_ms'], 'parity', d['parity']['ok'], d['clocks']['sm_mhz'])" | tee -a gpurun_out/quick_ab.log
done
