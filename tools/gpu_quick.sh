#!/bin/bash
# short GPU visit: the tests named in $TESTS (default: kernels + chain + step), then bench A/B over the env toggles in $AB
mkdir -p gpurun_out
TESTS=${TESTS:-"tests/test_kernels_gpu.py tests/test_conv2d_chain_gpu.py tests/test_step_gpu.py"}
timeout 900 python -m pytest $TESTS -m gpu -q --tb=short -x > gpurun_out/quick_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/quick_tests.log
tail -15 gpurun_out/quick_tests.log
run() { # name, env...
  local name=$1; shift
  env "$@" timeout 300 python bench.py --steps ${STEPS:-200} --warmup 20 --no-cpu-baseline --no-extras > gpurun_out/ab_$name.json 2> gpurun_out/ab_$name.err
  python - <<P
import json
try:
    d=json.load(open('gpurun_out/ab_$name.json'))
    print('%-28s value %.0f  e2e %.0f  critic %.3f ms  gen %.3f ms  launches c/g %s/%s' % ('$name', d['value'], d['e2e']['value'], d['critic_step_ms'], d['generator_step_ms'], d.get('kernels_per_critic_step'), d.get('kernels_per_generator_step')))
except Exception as e:
    print('$name failed', e); print(open('gpurun_out/ab_$name.err').read()[-1500:])
P
}
for spec in "${@}"; do
  name=${spec%%:*}; envs=${spec#*:}
  run $name ${envs//,/ }
done
