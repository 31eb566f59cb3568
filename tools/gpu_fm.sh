#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest.log
tail -12 gpurun_out/pytest.log
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_c3_b.json 2> gpurun_out/bench_c3_b.err; python -c "
import json; d=json.load(open('gpurun_out/bench_c3_b.json')); print('c3', d['value'], d['roofline']['achieved'], d['roofline']['fwd_kernel']['achieved'])"
timeout 600 python bench.py --workload c4 --steps 3 --warmup 2 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; tail -3 gpurun_out/bench_c4.err; python -c "
import json; d=json.load(open('gpurun_out/bench_c4.json')); print('c4', d['encodings'])"
