#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/ab3.log
timeout 1200 python -m pytest tests/test_flash_attn_gpu.py tests/test_features_gpu.py -m gpu -x -q 2>&1 | tail -2 | tee gpurun_out/ab3_pytest.log
D=$PWD/xhy-flash-attention_b200/lib_ab
for i in 1 2; do
  for v in $D/*.so ""; do FA_B200_LIB=$v python tools/len_sweep.py 2>&1 | tail -1 | tee -a gpurun_out/ab3.log; done
done
for v in $D/*.so ""; do FA_B200_LIB=$v python tools/ab_kernels.py 2>&1 | tee -a gpurun_out/ab3.log; done
for v in $D/*.so ""; do
  FA_B200_LIB=$v python bench.py --workload c4 --steps 5 --warmup 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('c4', '$v'[-12:] or 'new', {k:(round(v['tflops']), round(v['fwd_kernel_ms'],2), round(v['bwd_kernel_ms'],2)) for k,v in d['encodings'].items()})" | tee -a gpurun_out/ab3.log
done
