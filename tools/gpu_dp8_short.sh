#!/bin/bash
# short 8-GPU visit: weak scaling at 8 ranks, fp32 and bf16 (the smaller world sizes: tools/gpu_dp4.sh)
mkdir -p gpurun_out
B="bench.py --steps ${STEPS:-200} --warmup 20 --no-cpu-baseline --no-extras"
for prec in fp32 bf16; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29908 $B --gpus 8 --precision $prec > gpurun_out/dp8s_${prec}_n8.json 2> gpurun_out/dp8s_${prec}_n8.err
  python - <<P
import json
try:
    d = json.loads([l for l in open('gpurun_out/dp8s_${prec}_n8.json') if l.startswith('{')][-1])
    print('$prec N=8 value %.0f e2e %.0f ms/step %.3f critic %.3f pair %.3f' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['critic_step_ms'], d['critic_plus_generator_step_ms']))
except Exception as e:
    print('$prec failed', e); print(open('gpurun_out/dp8s_${prec}_n8.err').read()[-600:])
P
done
