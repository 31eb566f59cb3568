#!/bin/bash
# pass E: streamed-chunk softmax (sm100_fwd.cu) -- parity, timeline, A/B against the previous build; fwd2 OFULL fix on d=224
mkdir -p gpurun_out
P=$PWD/xhy-flash-attention_b200/lib_ab
for c in "fwd 64" "fwd 128" "bwd 128" "fwd 96" "fwd 256" ; do timeout 120 python tools/quick_check.py $c 2>&1 | grep -v Warning | tail -9; done > gpurun_out/quick.log 2>&1
grep -E "QUICK|!!" gpurun_out/quick.log
for d in 224 256; do timeout 120 python tools/repro_case.py 2 256 512 6 2 $d 0 bf16 8 2>&1 | tail -2; timeout 120 python tools/repro_case.py 2 1024 1024 6 2 $d 0 fp16 8 2>&1 | tail -2; done > gpurun_out/repro.log 2>&1
cat gpurun_out/repro.log
timeout 900 python -m pytest tests/test_flash_attn_gpu.py tests/test_features_gpu.py tests/test_fullsize_gpu.py -m gpu -q --timeout 600 > gpurun_out/pytest_fwd.log 2>&1; echo "pytest fwd exit $?"
grep -E "^FAILED|passed|failed" gpurun_out/pytest_fwd.log | head -20
FA_B200_LIB=$P/lib_trace.so timeout 120 python tools/trace_fwd.py > gpurun_out/trace_new.txt 2>&1
sed -n 1,24p gpurun_out/trace_new.txt
timeout 300 python tools/fwd_time.py streamed > gpurun_out/fwd_time_new.json 2> gpurun_out/fwd_time_new.err
FA_B200_LIB=$P/lib_base.so FA_B200_FWD2=0 timeout 300 python tools/fwd_time.py base > gpurun_out/fwd_time_old.json 2> gpurun_out/fwd_time_old.err
timeout 300 python tools/fwd_time.py streamed_again > gpurun_out/fwd_time_new2.json 2>> gpurun_out/fwd_time_new.err
python - <<'PY'
import json
for f in ('new','old','new2'):
    try:
        d=json.load(open(f'gpurun_out/fwd_time_{f}.json'))
    except Exception as e:
        print(f, 'unreadable', e); continue
    print(d.get('tag'), {k: round(v['fwd_tflops']) for k, v in d.items() if isinstance(v, dict) and 'fwd_tflops' in v})
PY
