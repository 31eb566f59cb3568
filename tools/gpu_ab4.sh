#!/bin/bash
mkdir -p gpurun_out
D=xhy-flash-attention_b200/lib_ab
AB_CASES="fmdoc,varlen" python tools/ab_inproc.py $D/lib_head.so new 2>&1 | tee gpurun_out/ab4.log
