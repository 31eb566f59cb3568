#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_decode_capi_gpu.py -m gpu -x -q > gpurun_out/pytest_decode.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_decode.log
timeout 300 python bench.py --workload c5b > gpurun_out/bench_c5b.json 2> gpurun_out/bench_c5b.err; echo "bench exit $?"
timeout 300 python tools/ncu_decode.py > gpurun_out/plain_decode.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'decode_kernel|combine_kernel' -s 2 -c 4 -o gpurun_out/prof_decode_r1 \
    python tools/ncu_decode.py > gpurun_out/ncu_decode.log 2>&1
tail -15 gpurun_out/pytest_decode.log; cat gpurun_out/bench_c5b.json; tail -3 gpurun_out/bench_c5b.err
