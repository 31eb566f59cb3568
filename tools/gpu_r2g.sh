#!/bin/bash
# pass G: backward with dQ first / dK from shared-memory dS^T: parity, timeline, timing against the previous build
mkdir -p gpurun_out
P=$PWD/xhy-flash-attention_b200/lib_ab
for c in "bwd 128" "bwd 64" "bwd 96"; do timeout 120 python tools/quick_check.py $c 2>&1 | grep -v Warning | tail -9; done > gpurun_out/quick.log 2>&1
grep -E "QUICK|!!" gpurun_out/quick.log
timeout 900 python -m pytest tests/test_flash_attn_gpu.py tests/test_features_gpu.py tests/test_fullsize_gpu.py tests/test_dropout_gpu.py tests/test_decode_capi_gpu.py tests/test_flashmask_capi_gpu.py -m gpu -q --timeout 600 -x > gpurun_out/pytest_bwd.log 2>&1; echo "pytest exit $?"
grep -E "^FAILED|passed|failed" gpurun_out/pytest_bwd.log | head -20
FA_B200_LIB=$P/lib_trace.so timeout 120 python tools/trace_bwd.py > gpurun_out/trace_bwd_new.txt 2>&1
sed -n 1,12p gpurun_out/trace_bwd_new.txt | cut -c1-220
timeout 300 python tools/fwd_time.py new > gpurun_out/fwd_time_new.json 2> gpurun_out/fwd_time_new.err
FA_B200_LIB=$P/lib_base.so FA_B200_FWD2=0 timeout 300 python tools/fwd_time.py base > gpurun_out/fwd_time_old.json 2> gpurun_out/fwd_time_old.err
timeout 300 python tools/fwd_time.py new_again > gpurun_out/fwd_time_new2.json 2>> gpurun_out/fwd_time_new.err
python - <<'PY'
import json
for f in ('new','old','new2'):
    try:
        d=json.load(open(f'gpurun_out/fwd_time_{f}.json'))
    except Exception as e:
        print(f, 'unreadable', e); continue
    print(d.get('tag'), {k[:14]: (round(v['fwd_tflops']), round(v['bwd_tflops'] or 0)) for k, v in d.items() if isinstance(v, dict) and 'fwd_tflops' in v})
PY
