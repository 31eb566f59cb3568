#!/bin/bash
for m in "FA_B200_FWD2=0" "FA_B200_FWD2=1" "FA_B200_FWD2=1 FA_B200_FWD2_QT=1"; do
  env $m timeout 300 python tools/fwd_time.py "$m" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['tag'], {k.split()[0]+' '+k.split()[-2]+('c' if 'causal' in k and 'non' not in k else 'n'): round(v.get('fwd_tflops',0)) for k,v in d.items() if isinstance(v,dict)})
"
done
