#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_features_gpu.py tests/test_decode_capi_gpu.py -m gpu -x -q > gpurun_out/pytest_feat.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_feat.log
tail -5 gpurun_out/pytest_feat.log
timeout 600 python bench.py --workload c4 --steps 3 --warmup 2 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; tail -3 gpurun_out/bench_c4.err; python -c "
import json; d=json.load(open('gpurun_out/bench_c4.json')); print('c4', d['encodings'])"
