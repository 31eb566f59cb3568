#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/ab.log
timeout 900 python -m pytest tests/test_flash_attn_gpu.py tests/test_features_gpu.py -m gpu -x -q 2>&1 | tail -3
D=$PWD/xhy-flash-attention_b200/lib_ab
for i in 1 2; do
  for v in $D/*.so; do FA_B200_LIB=$v python tools/ab_kernels.py 2>&1 | tee -a gpurun_out/ab.log; done
  python tools/ab_kernels.py 2>&1 | tee -a gpurun_out/ab.log
done
