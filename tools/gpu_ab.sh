#!/bin/bash
# in-process A/B of the in-tree library against variant builds under lib_ab/:  gpu_ab.sh "cases" lib_x.so [lib_y.so ...]
CASES=$1; shift
L=""; for x in "$@"; do L="$L xhy-flash-attention_b200/lib_ab/$x"; done
for x in "$@"; do FA_B200_LIB=xhy-flash-attention_b200/lib_ab/$x timeout 300 python tools/quick_check.py bwd 128 2>&1 | tail -2; done
AB_CASES="$CASES" AB_ROUNDS=10 timeout 900 python tools/ab_inproc.py new $L 2>&1 | grep -v Warning | tail -30
