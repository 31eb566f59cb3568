#!/bin/bash
mkdir -p gpurun_out
python tools/fm_probe.py ncu > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'fwd_kernel' -s 17 -c 1 -o gpurun_out/prof_r1_fm \
    python tools/fm_probe.py ncu > gpurun_out/ncu_fm.log 2>&1
tail -2 gpurun_out/ncu_fm.log
