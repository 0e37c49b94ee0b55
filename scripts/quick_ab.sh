#!/bin/bash
# GPU box: extraction + prep parity files, then the stage clocks of the default build (twice) and the single-frame latency.
mkdir -p gpurun_out
python -m pytest tests/test_gpu_extract.py tests/test_gpu_prep.py -x -q 2>&1 | tail -2 | tee gpurun_out/quick_ab.log
for i in 1 2; do
  python bench.py --steps 50 --warmup 3 --no-cpu-baseline --no-match 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('value', round(d['value']), 'e2e', round(d['e2e']['value']), d['roofline']['stage_ms'], 'parity', d['parity']['ok'], d['clocks']['sm_mhz'])" | tee -a gpurun_out/quick_ab.log
done
python scripts/bench_latency.py 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('latency', {k: round(v['gpu_ms_median'],4) for k,v in d['results'].items()})" | tee -a gpurun_out/quick_ab.log
