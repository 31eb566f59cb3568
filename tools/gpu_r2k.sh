#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_fp8_gpu.py -m gpu -q --timeout 120 > gpurun_out/pytest_fp8.log 2>&1; echo "pytest fp8 exit $?"
grep -E "^FAILED|^ERROR|passed|failed" gpurun_out/pytest_fp8.log | head -40
grep -n "Error\|assert \|^E " gpurun_out/pytest_fp8.log | head -30
timeout 300 python tools/fp8_time.py 2>&1 | tail -5
