#!/bin/bash
# multi-GPU visit (gpurun --gpus N): data-parallel parity tests at every world size up to N, then bench lines per world size
# usage: tools/gpu_dp.sh "<world sizes>" [bench variants name:ENV=..,ENV=.. ...]
mkdir -p gpurun_out
WORLDS=${1:-2}; shift
nvidia-smi --query-gpu=index,name --format=csv,noheader > gpurun_out/dp_gpus.txt
for w in $WORLDS; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $w --master-addr 127.0.0.1 --master-port $((29600 + w)) tools/dp_gpu_check.py > gpurun_out/dp_check_n$w.log 2>&1
  echo "dp_gpu_check world $w rc=$?"; grep -h "dp$w:\|ALL OK\|FAILED\|fused\|soak" gpurun_out/dp_check_n$w.log | tail -12
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29650 tools/dp_api_check.py > gpurun_out/dp_api_n2.log 2>&1; echo "dp_api_check rc=$?"; tail -2 gpurun_out/dp_api_n2.log
for spec in "$@"; do
  name=${spec%%:*}; envs=${spec#*:}
  for w in $WORLDS 1; do
    if [ "$w" == "1" ]; then
      env ${envs//,/ } timeout 600 python bench.py --gpus 1 --steps ${STEPS:-200} --warmup 20 --no-cpu-baseline --no-extras $BENCHARGS > gpurun_out/dpbench_${name}_n$w.json 2> gpurun_out/dpbench_${name}_n$w.err
    else
      env ${envs//,/ } timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $w --master-addr 127.0.0.1 --master-port $((29700 + w)) bench.py --gpus $w --steps ${STEPS:-200} --warmup 20 --no-cpu-baseline --no-extras $BENCHARGS > gpurun_out/dpbench_${name}_n$w.json 2> gpurun_out/dpbench_${name}_n$w.err
    fi
    python - <<P
import json
try:
    d=json.loads([l for l in open('gpurun_out/dpbench_${name}_n$w.json') if l.startswith('{')][-1])
    print('%-14s N=%d value %.0f  e2e %.0f  ms/step %.3f critic %.3f gen %.3f dtype %s' % ('$name', d['n_gpus'], d['value'], d['e2e']['value'], d['ms_per_step'], d['critic_step_ms'], d['generator_step_ms'], d['dtype']))
except Exception as e:
    print('$name N=$w failed', e); print(open('gpurun_out/dpbench_${name}_n$w.err').read()[-1500:])
P
  done
done
