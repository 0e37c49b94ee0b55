T=r02fin
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/${T}_pytest.log
python __graft_entry__.py smoke > gpurun_out/${T}_smoke.log 2>&1
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_ref.json 2>/dev/null
python scripts/bench_latency.py > gpurun_out/${T}_latency.json 2>/dev/null
cat gpurun_out/${T}_pytest.log; tail -1 gpurun_out/${T}_smoke.log; cat gpurun_out/${T}_bench.json | cut -c1-400
