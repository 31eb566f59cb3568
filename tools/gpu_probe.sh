#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/fm_probe.py all > gpurun_out/fm_probe.log 2>&1; cat gpurun_out/fm_probe.log
timeout 600 python tools/fm_probe.py ncu > gpurun_out/plain_fm.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'fwd_kernel' -s 20 -c 1 -o gpurun_out/prof_fm \
    python tools/fm_probe.py ncu > gpurun_out/ncu_fm.log 2>&1
tail -3 gpurun_out/ncu_fm.log
