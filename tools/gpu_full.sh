#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest.log
tail -15 gpurun_out/pytest.log
timeout 300 python bench.py --workload c5a --ring-tokens 32768 --steps 5 --warmup 3 > gpurun_out/bench_ring1.json 2> gpurun_out/bench_ring1.err; tail -3 gpurun_out/bench_ring1.err; cat gpurun_out/bench_ring1.json
