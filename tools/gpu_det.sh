#!/bin/bash
mkdir -p gpurun_out
D=xhy-flash-attention_b200/lib_ab
timeout 600 python -m pytest tests/test_flash_attn_gpu.py -m gpu -x -q -k "deterministic or race or cuda_graph" 2>&1 | tail -4 | tee gpurun_out/det_pytest.log
timeout 900 python -m pytest tests/test_flash_attn_gpu.py tests/test_features_gpu.py tests/test_decode_capi_gpu.py -m gpu -x -q 2>&1 | tail -2 | tee -a gpurun_out/det_pytest.log
AB_CASES="c3,s1024" timeout 600 python tools/ab_inproc.py $D/lib_head.so new 2>&1 | tee gpurun_out/ab_det.log
