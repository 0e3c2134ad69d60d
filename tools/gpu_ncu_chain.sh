#!/bin/bash
mkdir -p gpurun_out
timeout 120 python tools/chain_trace.py fwd > gpurun_out/plain_chain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:conv2d_chain -s 3 -c 1 -f -o gpurun_out/prof_chain python tools/chain_trace.py fwd > gpurun_out/ncu_chain.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_chain.log; ls -la gpurun_out/prof_chain.ncu-rep
