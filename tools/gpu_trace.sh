#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_conv2d_chain_gpu.py -q --tb=short -x 2>&1 | tail -3 > gpurun_out/chain_trace.log
for m in fwd dgrad tangent; do timeout 120 python tools/chain_trace.py $m; done >> gpurun_out/chain_trace.log 2>&1
N=30 timeout 120 python tools/chain_trace.py fwd 2>&1 | head -3 >> gpurun_out/chain_trace.log
PCTTS_CHAIN_PARTS=2 timeout 120 python tools/chain_trace.py fwd 2>&1 | head -12 >> gpurun_out/chain_trace.log
cat gpurun_out/chain_trace.log
