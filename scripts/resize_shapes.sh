#!/bin/bash
# GPU box: stage clocks for k_resize_fast CTA shapes (ORBX_RESIZE_SHAPE=bx,by) and k_blur strip heights (ORBX_BLUR_ROWS)
mkdir -p gpurun_out; rm -f gpurun_out/resize_shapes.log
line() { python bench.py --steps 40 --warmup 3 --no-cpu-baseline --no-match 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('$1 value', round(d['value']), d['roofline']['stage_ms'], 'parity', d['parity']['ok'])" | tee -a gpurun_out/resize_shapes.log; }
for v in ${SHAPES:-16,8 16,4 16,2 32,4}; do ORBX_RESIZE_SHAPE=$v line "shape $v"; done
for v in ${ROWS:-24 48 72 96 144}; do ORBX_BLUR_ROWS=$v line "blur_rows $v"; done
ORBX_BLUR_ROWS=96 python -m pytest tests/test_gpu_extract.py -x -q 2>&1 | tail -2 | tee -a gpurun_out/resize_shapes.log
