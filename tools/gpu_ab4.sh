#!/bin/bash
mkdir -p gpurun_out
D=xhy-flash-attention_b200/lib_ab
timeout 1200 python -m pytest tests/test_flash_attn_gpu.py tests/test_features_gpu.py -m gpu -x -q 2>&1 | tail -2 | tee gpurun_out/ab4_pytest.log
python tools/ab_inproc.py $D/lib_head.so new 2>&1 | tee gpurun_out/ab4.log
