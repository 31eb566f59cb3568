#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dropout_gpu.py -m gpu -q --timeout 300 -x 2>&1 | tail -3
timeout 600 python - <<'PY' 2>&1 | grep -v Warning | tail -5
import json, bench
cx = bench.Ctx()
print(json.dumps(bench.dropout_measure(cx)))
PY
