#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest.log
tail -4 gpurun_out/pytest.log
for w in c3 c2; do
timeout 300 python bench.py --workload $w --steps 20 --warmup 5 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; python -c "
import json; d=json.load(open('gpurun_out/bench_$w.json')); print('$w', round(d['value'],1), 'ms', round(d['ms_per_step'],3), 'bwd', round(d['roofline']['achieved'],1), d['roofline']['ms_per_launch'], 'fwd', round(d['roofline']['fwd_kernel']['achieved'],1), d['roofline']['fwd_kernel']['ms_per_launch'], d['clocks'])"
done
timeout 600 python bench.py --workload c4 --steps 3 --warmup 2 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; tail -3 gpurun_out/bench_c4.err; python -c "
import json; d=json.load(open('gpurun_out/bench_c4.json')); print('c4', d['encodings'])"
