#!/bin/bash
# First-contact GPU diagnostics: every test module in its own process (a trapped kernel poisons the CUDA
# context of its process only), bounded by timeout so a protocol bug cannot hang the box.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/diag_gpu.txt 2>&1
for sel in "tests/test_gpu_ops.py -k linear" "tests/test_gpu_ops.py -k conv1d" "tests/test_gpu_ops.py -k attention" \
           "tests/test_gpu_forward.py" "tests/test_gpu_inference.py"; do
  name=$(echo "$sel" | tr ' /.' '___')
  echo "=== $sel" | tee -a gpurun_out/diag_summary.txt
  timeout 600 python -m pytest $sel -m gpu -q --tb=short -rA -p no:cacheprovider > gpurun_out/diag_$name.log 2>&1
  echo "exit=$?" | tee -a gpurun_out/diag_summary.txt
  tail -n 3 gpurun_out/diag_$name.log | tee -a gpurun_out/diag_summary.txt
done
