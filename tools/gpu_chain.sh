#!/bin/bash
# GPU visit for the chain kernel: its own parity tests first (bounded by timeout: a hang must not eat the box), then the step tests
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_conv2d_chain_gpu.py -q --tb=short -x > gpurun_out/chain_tests.log 2>&1; echo "chain rc=$?" >> gpurun_out/chain_tests.log
tail -30 gpurun_out/chain_tests.log
