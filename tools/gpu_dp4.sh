#!/bin/bash
# 4-GPU visit: weak scaling at 4 / 2 / 1 ranks (fp32) and bf16 at 4 / 1, the smaller runs sharing the box on disjoint GPU sets
mkdir -p gpurun_out
STEPS=${STEPS:-200}
B="bench.py --steps $STEPS --warmup 20 --no-cpu-baseline --no-extras"
tr() { local gpus=$1 world=$2 port=$3 out=$4; shift 4
  if [ "$world" == "1" ]; then CUDA_VISIBLE_DEVICES=$gpus timeout 600 python "$@" > gpurun_out/$out.json 2> gpurun_out/$out.err
  else CUDA_VISIBLE_DEVICES=$gpus timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $world --master-addr 127.0.0.1 --master-port $port "$@" > gpurun_out/$out.json 2> gpurun_out/$out.err; fi; }
tr 0,1,2,3 4 29804 dp4_fp32_n4 $B --gpus 4
tr 0,1,2,3 4 29814 dp4_bf16_n4 $B --gpus 4 --precision bf16
tr 0,1 2 29802 dp4_fp32_n2 $B --gpus 2 & tr 2 1 0 dp4_fp32_n1 $B --gpus 1 & tr 3 1 0 dp4_bf16_n1 $B --gpus 1 --precision bf16 & wait
python - <<'P'
import json, glob, os
for f in sorted(glob.glob('gpurun_out/dp4_*_n*.json')):
    try:
        d = json.loads([l for l in open(f) if l.startswith('{')][-1])
        print('%-14s N=%d value %9.0f e2e %9.0f ms/step %.3f critic %.3f pair %.3f %s' % (os.path.basename(f)[4:-5], d['n_gpus'], d['value'], d['e2e']['value'], d['ms_per_step'], d['critic_step_ms'], d['critic_plus_generator_step_ms'], d['dtype']))
    except Exception as e:
        print(f, 'FAILED', e); print(open(f[:-5] + '.err').read()[-500:])
P
