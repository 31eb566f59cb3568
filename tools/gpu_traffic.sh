#!/bin/bash
# DRAM traffic of the C3 kernels (b8 s8192 h32/8 d128 causal): plain run first, then one ncu --set full of one fwd + one bwd launch
mkdir -p gpurun_out
python tools/ncu_case.py 8 8192 32 8 1 2 > gpurun_out/plain_case_c3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'fwd_kernel|bwd_kernel' -s 2 -c 2 -o gpurun_out/prof_r1_c3 \
    python tools/ncu_case.py 8 8192 32 8 1 2 > gpurun_out/ncu_full_c3.log 2>&1
tail -2 gpurun_out/ncu_full_c3.log
timeout 900 python -m pytest tests/test_features_gpu.py -m gpu -q -k "packed_variants" 2>&1 | tail -2
