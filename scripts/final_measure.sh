#!/bin/bash
# GPU box: the round's record — GPU tests, smoke, every BASELINE config as a bench line, reference arm, launch list.
T=${1:-r02z}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/${T}_pytest.log
python __graft_entry__.py smoke > gpurun_out/${T}_smoke.log 2>&1
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
for c in 2 3 4 5; do python bench.py --config $c > gpurun_out/${T}_config$c.json 2> gpurun_out/${T}_config$c.err; done
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_ref.json 2>/dev/null
python scripts/bench_latency.py > gpurun_out/${T}_latency.json 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-match > gpurun_out/${T}_ncu_bench.log 2>&1
cat gpurun_out/${T}_pytest.log; tail -1 gpurun_out/${T}_smoke.log
# one full-set capture of the dominant kernel (launch 10 = a 192-frame batch launch; 0-7 are the per-level launches of the parity gate), summarised on the build side into profiles/
ncu --set full --clock-control none --import-source on -k regex:k_fast_t -s 10 -c 1 -f -o gpurun_out/${T}_k_fast python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-match > gpurun_out/${T}_ncu_fast.log 2>&1
