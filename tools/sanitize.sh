#!/bin/bash
# compute-sanitizer over smoke() and three operator tests: ONE tool per gpurun call.
#   bash tools/sanitize.sh memcheck|racecheck|synccheck|initcheck
# The library is rebuilt with a 100x longer mbarrier wait bound (the sanitizer slows kernels that much; the default
# bound would trap), run, and rebuilt as shipped.  Log: gpurun_out/sanitize_<tool>.log
set -u
cd "$(dirname "$0")/.."
TOOL=${1:-memcheck}
mkdir -p gpurun_out
TTS_EXTRA_NVCC_FLAGS="-DTTS_WAIT_BOUND_CYCLES=400000000000LL" python -m transformertts_b200.build --force > /dev/null || exit 1
LOG=gpurun_out/sanitize_$TOOL.log
: > $LOG
run() {
  echo "=== compute-sanitizer --tool $TOOL $*" | tee -a $LOG
  timeout 900 compute-sanitizer --tool $TOOL --print-limit 20 --error-exitcode 9 "$@" >> $LOG 2>&1
  echo "exit=$?" | tee -a $LOG
}
run python __graft_entry__.py smoke
run python -m pytest tests/test_gpu_ops.py -m gpu -q -x -p no:cacheprovider -k "test_linear_residual_layernorm or test_ffn_block[333] or test_attention_cross_with_probabilities[2-200-150] or test_attention_causal_with_padding[2-300] or test_conv1d[3-150-256-256]"
run python -m pytest tests/test_gpu_train_aux.py -m gpu -q -x -p no:cacheprovider -k "test_mix_matches_reference_golden or test_loss_matches_reference_golden"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|exit=|passed|failed" $LOG | tail -n 20
python -m transformertts_b200.build --force > /dev/null
