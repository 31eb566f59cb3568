#!/bin/bash
# One GPU-box pass: parity tests, smoke, every bench workload + reference arm, ncu launch list, ncu full of the top kernels.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/smoke.log
python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench exit $?"
python bench.py --workload c2 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err
python bench.py --workload c4 --steps 5 --warmup 3 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err
python bench.py --workload c5b > gpurun_out/bench_c5b.json 2> gpurun_out/bench_c5b.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
python bench.py --steps 2 --warmup 3 > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
python tools/ncu_case.py 8 8192 32 8 1 2 > gpurun_out/plain_case.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'fwd_kernel|bwd_kernel' -s 2 -c 2 -o gpurun_out/prof_r1 \
    python tools/ncu_case.py 8 8192 32 8 1 2 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/pytest.log; tail -2 gpurun_out/smoke.log; for w in c3 c2 c4 c5b; do python -c "
import json; d=json.load(open('gpurun_out/bench_$w.json')); print('$w', round(d['value'],1), d['unit'], round(d['ms_per_step'],3), 'ms', d.get('e2e',{}).get('value'), d.get('roofline',{}).get('frac'), d.get('encodings'))"; done
