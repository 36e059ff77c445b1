#!/bin/bash
# compute-sanitizer over the small-shape GPU tests (hand-rolled mbarrier / TMEM / TMA pipelines, nine epilogue modes):
#   memcheck  -- out-of-bounds / misaligned global and shared accesses
#   synccheck -- invalid barrier usage
#   racecheck -- shared-memory hazards between warps of a CTA
# Usage (on the GPU box): bash tools/gpu_sanitize.sh [outdir]     Logs: <outdir>/sanitize_<tool>.log (+ a summary file)
# Each tool runs under its own timeout; the selection keeps one case per kernel family (sanitized kernels run 10-100x slower).
set -u
OUT=${1:-gpurun_out}
mkdir -p "$OUT"
SEL='test_conv_store_matches_oracle or test_conv_pool_and_argmax or test_conv_head or test_ctx_pool_matches_oracle or test_mse_count or test_sgd_matches or test_unpool_routes or test_first_layer_fused or (test_forward_matches_reference_golden and he_1x64x96) or (test_parameter_gradients_match_oracle and case0 and bf16)'
SUMMARY="$OUT/sanitize_summary.txt"
: > "$SUMMARY"
for TOOL in memcheck synccheck racecheck; do
  LOG="$OUT/sanitize_${TOOL}.log"
  timeout ${SANITIZE_TIMEOUT:-900} compute-sanitizer --tool "$TOOL" --error-exitcode 86 --print-limit 20 \
      python -m pytest tests/test_gpu_forward.py tests/test_gpu_backward.py -q -x -m gpu -k "$SEL" -p no:cacheprovider \
      > "$LOG" 2>&1
  RC=$?
  {
    echo "== $TOOL: exit code $RC (0 = tests passed and no finding, 86 = sanitizer finding, 124 = timeout)"
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" "$LOG" | tail -4
  } >> "$SUMMARY"
done
cat "$SUMMARY"
