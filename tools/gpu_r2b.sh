#!/bin/bash
# round-2 pass B: split-row forward (sm100_fwd2.cu) + sliced wide backward (sm100_bwd_wide.cu): quick parity per kernel in
# separate processes, then the whole suite, then A/B timing against the round-1 forward, then the default bench line
mkdir -p gpurun_out
for c in "fwd 64" "fwd 128" "fwd 192" "fwd 256" "fwd 96" "bwd 256" "bwd 192" "bwd 160" "bwd 128"; do
  timeout 120 python tools/quick_check.py $c 2>&1 | grep -v Warning | tail -9
done > gpurun_out/quick.log 2>&1
cat gpurun_out/quick.log
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 -x > gpurun_out/pytest.log 2>&1; echo "pytest all exit $?"
tail -30 gpurun_out/pytest.log
FA_B200_FWD2=1 timeout 300 python tools/fwd_time.py new > gpurun_out/fwd_time_new.json 2> gpurun_out/fwd_time_new.err; echo "time new exit $?"
FA_B200_FWD2=0 timeout 300 python tools/fwd_time.py old > gpurun_out/fwd_time_old.json 2> gpurun_out/fwd_time_old.err; echo "time old exit $?"
FA_B200_FWD2=1 timeout 300 python tools/fwd_time.py new2 > gpurun_out/fwd_time_new2.json 2>> gpurun_out/fwd_time_new.err
cat gpurun_out/fwd_time_new.json gpurun_out/fwd_time_old.json gpurun_out/fwd_time_new2.json
timeout 900 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench exit $?"
tail -5 gpurun_out/bench_c3.err
python - <<'P'
import json
d=json.load(open('gpurun_out/bench_c3.json'))
print('c3', round(d['value'],1), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), 'roof', round(d['roofline']['frac'],3), 'fwd', round(d['roofline']['fwd_kernel']['achieved'],1))
for k in ('workloads','yardstick'):
    print(k, json.dumps(d.get(k))[:3500])
P
