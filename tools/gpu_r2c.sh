#!/bin/bash
# round-2 pass C: flaky-case probe, whole suite without -x, forward A/B (old | fwd2 two tiles | fwd2 one tile)
mkdir -p gpurun_out
for i in 1 2; do timeout 120 python tools/repro_case.py 2 108 256 6 2 224 0 fp16 12 2>&1 | tail -4; done > gpurun_out/repro.log 2>&1
timeout 120 python tools/repro_case.py 2 108 256 6 2 224 0 bf16 12 2>&1 | tail -3 >> gpurun_out/repro.log
timeout 120 python tools/repro_case.py 2 108 256 6 2 256 0 fp16 12 2>&1 | tail -3 >> gpurun_out/repro.log
timeout 120 python tools/repro_case.py 2 108 256 6 2 192 0 fp16 12 2>&1 | tail -3 >> gpurun_out/repro.log
cat gpurun_out/repro.log
timeout 1800 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/pytest.log 2>&1; echo "pytest all exit $?"
grep -E "^FAILED|passed|failed" gpurun_out/pytest.log | head -40
FA_B200_FWD2=1 FA_B200_FWD2_QT=1 timeout 300 python tools/fwd_time.py one_tile > gpurun_out/fwd_time_qt1.json 2> gpurun_out/fwd_time_qt1.err; echo "time qt1 exit $?"
FA_B200_FWD2=0 timeout 300 python tools/fwd_time.py old > gpurun_out/fwd_time_old.json 2> gpurun_out/fwd_time_old.err
FA_B200_FWD2=1 timeout 300 python tools/fwd_time.py two_tiles > gpurun_out/fwd_time_new.json 2> gpurun_out/fwd_time_new.err
FA_B200_FWD2=1 FA_B200_FWD2_QT=1 timeout 300 python tools/fwd_time.py one_tile_again > gpurun_out/fwd_time_qt1b.json 2>> gpurun_out/fwd_time_qt1.err
python - <<'P'
import json
for f in ('qt1','old','new','qt1b'):
    try:
        d=json.load(open(f'gpurun_out/fwd_time_{f}.json'))
    except Exception as e:
        print(f, 'unreadable', e); continue
    print(d.get('tag'), {k: round(v['fwd_tflops']) for k, v in d.items() if isinstance(v, dict) and 'fwd_tflops' in v})
P
