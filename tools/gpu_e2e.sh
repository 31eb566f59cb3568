#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_flash_attn_gpu.py -m gpu -x -q -k "host_pipeline or qkvpacked" 2>&1 | tail -5
timeout 400 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; tail -5 gpurun_out/bench_c3.err; python -c "
import json; d=json.load(open('gpurun_out/bench_c3.json')); print('c3', round(d['value'],1), d['e2e'], d['gpu_launches'])"
