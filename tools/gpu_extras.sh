#!/bin/bash
# Decode-extras pass: the new parity tests, the decode bench (regression check) and the C3 bench (e2e pipeline).
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_kvcache_extras_gpu.py tests/test_handle_abi.py tests/test_decode_capi_gpu.py -m gpu -q \
    > gpurun_out/extras_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/extras_pytest.log
python bench.py --workload c5b > gpurun_out/x_bench_c5b.json 2> gpurun_out/x_bench_c5b.err
python bench.py > gpurun_out/x_bench_c3.json 2> gpurun_out/x_bench_c3.err
grep -E "passed|failed|FAILED|Error" gpurun_out/extras_pytest.log | tail -30
for w in c5b c3; do python -c "
import json; d=json.load(open('gpurun_out/x_bench_$w.json')); print('$w', round(d['value'],1), d['unit'], round(d['ms_per_step'],3), 'ms', d.get('e2e'))"; done
