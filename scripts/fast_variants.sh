#!/bin/bash
# GPU box: parity of the FAST kernel variants (whole extraction test file under each), then the stage clock of each.
mkdir -p gpurun_out
for v in ${FAST_TEST_VARIANTS:-1 3}; do ORBX_FAST=$v python -m pytest tests/test_gpu_extract.py -x -q 2>&1 | tail -2; done
for v in ${FAST_BENCH_VARIANTS:-"1 0" "3 0" "3 3" "3 5"}; do
  set -- $v
  ORBX_FAST=$1 ORBX_FAST3_NF=$2 python bench.py --steps 50 --warmup 3 --no-cpu-baseline --no-match 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('FAST=$1 NF=$2 value', round(d['value']), 'e2e', round(d['e2e']['value']), d['roofline']['stage_ms'], 'parity', d['parity']['ok'], d['clocks']['sm_mhz'])" | tee -a gpurun_out/fast_variants.log
done
