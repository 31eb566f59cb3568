#!/bin/bash
mkdir -p gpurun_out
python tools/ncu_case.py 2 8192 32 8 1 2 > gpurun_out/plain_case.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'fwd_kernel|bwd_kernel' -s 2 -c 2 -o gpurun_out/prof_r1b \
    python tools/ncu_case.py 2 8192 32 8 1 2 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
