#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/pytest.log 2>&1; echo "pytest all exit $?"
grep -E "^FAILED|^ERROR|passed|failed" gpurun_out/pytest.log | head -30
grep -n "Error\|assert" gpurun_out/pytest.log | head -20
