#!/bin/bash
# `ncu --set full` of the six kernel classes bench.py lists in `kernels` (one launch each, cfg-2 shapes), summaries under
# gpurun_out/ (copy to profiles/): raw metrics via tools/ncu_top.py and the per-source-line stall table via ncu_lines.py.
# ONE gpurun call; every shape first runs once without ncu.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for spec in "dec qkv|gemm2_kernel|gemm2_qkv" "dec out+LN|gemm2_kernel|gemm2_out_ln" "postnet conv 256->256|gemm2_kernel|gemm2_conv" \
            "dec ffn fused|ffn_kernel|ffn_fused" "dec self-attn|attn2_kernel|attn2_self" "dec cross-attn|attn3_kernel|attn3_cross" "dec cq|gemm2_kernel|gemm2_cross_q" "dec prenet2|gemm2_kernel|gemm2_prenet2" "dec prenet1|gemm2_kernel|gemm2_prenet1" "heads|gemm2_kernel|gemm2_heads" "postnet conv 80->256|gemm2_kernel|gemm2_conv_first" "postnet conv 256->80|gemm2_kernel|gemm2_conv_last"; do
  IFS='|' read -r flt kern key <<< "$spec"
  python tools/profile_op.py "$flt" > gpurun_out/plain_$key.log 2>&1 || { echo "plain run failed: $flt"; continue; }
  ncu --set full --clock-control none --import-source on -k regex:$kern -s 2 -c 1 -o gpurun_out/r02_$key \
      python tools/profile_op.py "$flt" > gpurun_out/ncu_$key.log 2>&1
  echo "$key: $(tail -1 gpurun_out/ncu_$key.log)"
done
