#!/bin/bash
# Round-2 evidence pass on one B200: parity suite, smoke, the driver's bench line, reference arm, ncu launch list of the
# headline run, one ncu --set full of the forward/backward kernels at head dims 128 and 256.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 1800 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest.log
grep -E "^FAILED|^ERROR|passed|failed" gpurun_out/pytest.log | head -20
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench exit $?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 300 python bench.py --no-extras --steps 2 --warmup 3 > gpurun_out/plain_bench.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --no-extras --steps 2 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
timeout 300 python tools/ncu_case.py 8 8192 32 8 1 2 128,256 > gpurun_out/plain_case.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'fwd_kernel|bwd_kernel|fwd2_kernel|bwd_wide_kernel' -c 8 -o gpurun_out/prof_r2 \
    python tools/ncu_case.py 8 8192 32 8 1 2 128,256 > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
python - <<PY
import json
d=json.load(open('gpurun_out/bench_c3.json'))
print('c3', round(d['value'],1), round(d['ms_per_step'],3), 'ms e2e', round(d['e2e']['value'],1), 'roof', round(d['roofline']['frac'],3), 'fwd', round(d['roofline']['fwd_kernel']['frac'],3), d['clocks'])
print(json.dumps(d.get('workloads'))[:1500])
print(json.dumps(d.get('yardstick'))[:1500])
PY
