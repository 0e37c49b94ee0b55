#!/bin/bash
# Parity tests + a short device/e2e bench on the GPU box (run through gpurun).
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-match "$@" | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('value', round(d['value']), 'e2e', round(d['e2e']['value']), d['roofline']['stage_ms'], d['clocks'])"
