#!/bin/bash
# the no-alloc entry point: alone in a fresh process, then the whole GPU suite (it is collected last)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 120 > gpurun_out/pytest_noalloc.log 2>&1; echo "alone exit $?"
tail -15 gpurun_out/pytest_noalloc.log | cut -c1-300
timeout 1800 python -m pytest tests -m gpu -x -q --timeout 600 > gpurun_out/pytest.log 2>&1; echo "pytest all exit $?"
grep -E "^FAILED|^ERROR|passed|failed" gpurun_out/pytest.log | head -20
