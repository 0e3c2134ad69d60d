#!/bin/bash
# ncu --set full of the named kernels at the demo shapes (tools/kernel_bench.py launches each one alone); one GPU, run after
# the plain command exited 0.  usage: tools/gpu_full_capture.sh <kernel_bench --only substring> <ncu -k regex> <out name>
mkdir -p gpurun_out
only="$1"; rex="$2"; out="${3:-full}"
python tools/kernel_bench.py --only "$only" --iters 2 > gpurun_out/${out}_plain.jsonl 2> gpurun_out/${out}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${out}_plain.err; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"$rex" -c 6 -o gpurun_out/${out} -f \
    python tools/kernel_bench.py --only "$only" --iters 2 > gpurun_out/${out}_ncu.log 2>&1
echo "ncu full ($out) rc=$?"; tail -2 gpurun_out/${out}_ncu.log
python tools/ncu_summary.py gpurun_out/${out}.ncu-rep > gpurun_out/${out}_summary.txt 2>&1
