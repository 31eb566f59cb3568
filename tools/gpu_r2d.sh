#!/bin/bash
# pass D: NaN-poison variant on head dims with partially out-of-bounds TMA boxes; forward timeline trace
mkdir -p gpurun_out
P=$PWD/xhy-flash-attention_b200/lib_ab
for d in 224 160 96 256; do
  FA_B200_LIB=$P/lib_poison.so FA_B200_FWD2=1 timeout 120 python tools/repro_case.py 2 256 512 6 2 $d 0 bf16 6 2>&1 | tail -3
done > gpurun_out/poison.log 2>&1
cat gpurun_out/poison.log
FA_B200_LIB=$P/lib_trace.so FA_B200_FWD2=0 timeout 120 python tools/trace_fwd.py > gpurun_out/trace_old.txt 2>&1
cat gpurun_out/trace_old.txt | head -50
