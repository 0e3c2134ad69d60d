#!/bin/bash
# One visit of an 8-GPU box (gpurun --gpus 8): data-parallel parity at 8 / 4 / 2 ranks, weak scaling fp32 and bf16 (BASELINE configs
# #2 / #3), strong scaling (global batch 8 over 1 / 2 / 4 / 8 ranks, global batch 10 over 1 / 2).  Runs that need fewer than 8 GPUs
# share the box on disjoint GPU sets (CUDA_VISIBLE_DEVICES), so the visit stays short.
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,clocks.max.sm --format=csv,noheader > gpurun_out/dp8_gpus.txt
STEPS=${STEPS:-200}
tr() {  # gpus(csv) world port script args... > out
  local gpus=$1 world=$2 port=$3 out=$4; shift 4
  if [ "$world" == "1" ]; then
    CUDA_VISIBLE_DEVICES=$gpus timeout 600 python "$@" > gpurun_out/$out.json 2> gpurun_out/$out.err
  else
    CUDA_VISIBLE_DEVICES=$gpus timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $world --master-addr 127.0.0.1 --master-port $port "$@" > gpurun_out/$out.json 2> gpurun_out/$out.err
  fi
}
B="bench.py --steps $STEPS --warmup 20 --no-cpu-baseline --no-extras"
echo "== parity"; 
tr 0,1,2,3,4,5,6,7 8 29608 dp8_check_n8 tools/dp_gpu_check.py; echo "check8 rc=$?"
tr 0,1,2,3 4 29604 dp8_check_n4 tools/dp_gpu_check.py & tr 4,5 2 29602 dp8_check_n2 tools/dp_gpu_check.py & tr 6,7 2 29612 dp8_api_n2 tools/dp_api_check.py & wait
for w in 8 4 2; do grep -h "dp$w:\|ALL OK\|FAILED\|fused all-reduce\|soak" gpurun_out/dp8_check_n$w.json | sort | uniq | head -8; done; tail -1 gpurun_out/dp8_api_n2.json
echo "== weak fp32"
tr 0,1,2,3,4,5,6,7 8 29708 dp8_fp32_n8 $B --gpus 8
tr 0,1,2,3 4 29704 dp8_fp32_n4 $B --gpus 4 & tr 4,5 2 29702 dp8_fp32_n2 $B --gpus 2 & tr 6 1 0 dp8_fp32_n1 $B --gpus 1 & tr 7 1 0 dp8_bf16_n1 $B --gpus 1 --precision bf16 & wait
echo "== weak bf16"
tr 0,1,2,3,4,5,6,7 8 29718 dp8_bf16_n8 $B --gpus 8 --precision bf16
tr 0,1,2,3 4 29714 dp8_bf16_n4 $B --gpus 4 --precision bf16 & tr 4,5 2 29712 dp8_bf16_n2 $B --gpus 2 --precision bf16 & tr 6,7 2 29722 dp8_strong10_n2 $B --gpus 2 --global-batch 10 & wait
echo "== strong"
tr 0,1,2,3,4,5,6,7 8 29728 dp8_strong8_n8 $B --gpus 8 --global-batch 8
tr 0,1,2,3 4 29724 dp8_strong8_n4 $B --gpus 4 --global-batch 8 & tr 4,5 2 29732 dp8_strong8_n2 $B --gpus 2 --global-batch 8 & tr 6 1 0 dp8_strong8_n1 $B --gpus 1 --global-batch 8 & tr 7 1 0 dp8_strong10_n1 $B --gpus 1 --global-batch 10 & wait
python - <<'P'
import json, glob, os
for f in sorted(glob.glob('gpurun_out/dp8_*_n*.json')):
    if 'check' in f or 'api' in f: continue
    try:
        d = json.loads([l for l in open(f) if l.startswith('{')][-1])
        print('%-22s N=%d B/GPU=%d value %9.0f e2e %9.0f ms/step %.3f critic %.3f gen %.3f %s %s' % (os.path.basename(f)[4:-5], d['n_gpus'], d['config']['global_batch_utterances'] // d['n_gpus'], d['value'], d['e2e']['value'], d['ms_per_step'], d['critic_step_ms'], d['generator_step_ms'], d['dtype'], d['scaling']))
    except Exception as e:
        print(f, 'FAILED', e); print(open(f[:-5] + '.err').read()[-600:])
P
