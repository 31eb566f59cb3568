#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dropout_gpu.py -m gpu -q --timeout 300 -x > gpurun_out/pytest_drop.log 2>&1; echo "pytest dropout exit $?"
grep -E "^FAILED|^ERROR|passed|failed" gpurun_out/pytest_drop.log | head -20
grep -n "Error\|assert " gpurun_out/pytest_drop.log | head -20
