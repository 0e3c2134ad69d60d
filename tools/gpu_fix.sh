#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_bn_sync_gpu.py tests/test_next_rows_gpu.py tests/test_precision_gpu.py tests/test_step_gpu.py -q --tb=short -s > gpurun_out/fix_tests.log 2>&1; echo "rc=$?" >> gpurun_out/fix_tests.log
grep -n "passed\|failed\|^FAILED\|gp input gradient\|precision " gpurun_out/fix_tests.log | tail -30
timeout 600 python tools/fullsize_diag.py > gpurun_out/fullsize_diag.log 2>&1; cat gpurun_out/fullsize_diag.log | tail -30
