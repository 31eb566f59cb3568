#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/ab.log
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
D=$PWD/xhy-flash-attention_b200/lib_ab
for i in 1 2; do
  FA_B200_LIB=$D/lib_e1412da.so python tools/ab_kernels.py 2>&1 | tee -a gpurun_out/ab.log
  python tools/ab_kernels.py 2>&1 | tee -a gpurun_out/ab.log
done
timeout 600 python bench.py --workload c4 --steps 3 --warmup 2 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; tail -3 gpurun_out/bench_c4.err; python -c "
import json; d=json.load(open('gpurun_out/bench_c4.json')); print('c4', d['encodings'])"
