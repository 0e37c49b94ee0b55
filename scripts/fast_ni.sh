#!/bin/bash
# GPU box: variant 1 of k_fast with NI max-by-IMAD pairs (ORBX_FAST1_NI): parity of the whole extraction file under the
# settings named in FAST_NI_TEST, then the stage clock of each setting in FAST_NI.
mkdir -p gpurun_out
for v in ${FAST_NI_TEST:-0 1}; do ORBX_FAST1_NI=$v python -m pytest tests/test_gpu_extract.py -x -q 2>&1 | tail -2 | tee -a gpurun_out/fast_ni.log; done
for v in ${FAST_NI:-0 1 2 5 0 1}; do
  ORBX_FAST1_NI=$v python bench.py --steps 50 --warmup 3 --no-cpu-baseline --no-match 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('NI=$v value', round(d['value']), 'e2e', round(d['e2e']['value']), d['roofline']['stage_ms'], 'parity', d['parity']['ok'], d['clocks']['sm_mhz'])" | tee -a gpurun_out/fast_ni.log
done
