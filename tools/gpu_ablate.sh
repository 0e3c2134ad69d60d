#!/bin/bash
# exposed time of each kernel family in the captured step graphs: bench with the family's C-ABI entry points stubbed out
mkdir -p gpurun_out
run() {
  local name=$1; shift
  env "$@" timeout 300 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-extras > gpurun_out/abl_$name.json 2> gpurun_out/abl_$name.err
  python - <<P
import json
try:
    d=json.load(open('gpurun_out/abl_$name.json'))
    print('%-22s critic %.3f ms  gen %.3f ms' % ('$name', d['critic_step_ms'], d['generator_step_ms']))
except Exception as e:
    print('$name failed', e); print(open('gpurun_out/abl_$name.err').read()[-800:])
P
}
run full X=1
run chain PCTTS_ABLATE=pctts_conv2d_chain
run wgrad_multi PCTTS_ABLATE=pctts_conv2d_wgrad_multi
run conv1d_fwd PCTTS_ABLATE=pctts_conv1d_tc_fwd
run conv1d_wgrad PCTTS_ABLATE=pctts_conv1d_tc_wgrad
run gemm_tc PCTTS_ABLATE=pctts_gemm_tc
run split PCTTS_ABLATE=pctts_split_bf16
run conv2d_tc PCTTS_ABLATE=pctts_conv2d_tc_fwd_fused
run bn PCTTS_ABLATE=pctts_bn_stats_partial,pctts_bn_finalize_partials,pctts_bn_apply,pctts_bn_apply_planes,pctts_bn_bwd_reduce_partial,pctts_bn_sync_bwd,pctts_bn_bwd_apply,pctts_bn_bwd_apply_planes
run lstm PCTTS_ABLATE=pctts_lstm_fwd,pctts_lstm_bwd
run colsum PCTTS_ABLATE=pctts_colsum
run small PCTTS_ABLATE=pctts_sum_scaled,pctts_fill_segments,pctts_loss_combine,pctts_gp_norms,pctts_gp_scale,pctts_uniform,pctts_critic_pack_inputs,pctts_add,pctts_copy2d,pctts_pad_time
run adam PCTTS_ABLATE=pctts_adam,pctts_split_bf16_multi
