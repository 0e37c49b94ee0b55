#!/bin/bash
# GPU box: stage clocks for k_resize_fast CTA shapes (ORBX_RESIZE_SHAPE=bx,by). (The k_blur strip heights in
# profiles/r02final_resize_blur_ab.log were measured with a run-time height that has since become a template parameter.)
mkdir -p gpurun_out; rm -f gpurun_out/resize_shapes.log
line() { python bench.py --steps 40 --warmup 3 --no-cpu-baseline --no-match 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('$1 value', round(d['value']), d['roofline']['stage_ms'], 'parity', d['parity']['ok'])" | tee -a gpurun_out/resize_shapes.log; }
for v in ${SHAPES:-16,8 16,4 16,2 32,4}; do ORBX_RESIZE_SHAPE=$v line "shape $v"; done
