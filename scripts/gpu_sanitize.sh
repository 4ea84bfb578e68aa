#!/bin/bash
# compute-sanitizer over the hot path: ONE tool per gpurun call (B200_PROFILING.md).  usage: bash scripts/gpu_sanitize.sh memcheck|initcheck|racecheck|synccheck
# Runs __graft_entry__.smoke() (the whole forward, N=512, 3 layers, through the C ABI) and a reduced set of kernel tests.
TOOL=${1:-memcheck}
mkdir -p gpurun_out
LOG=gpurun_out/sanitize_$TOOL.log
{
echo "== compute-sanitizer --tool $TOOL: smoke()"
timeout 900 compute-sanitizer --tool $TOOL --error-exitcode 99 python __graft_entry__.py --smoke 2>&1 | grep -v "^smoke: .* kernels" | tail -12
echo "exit code ${PIPESTATUS[0]}"
echo "== compute-sanitizer --tool $TOOL: kernel tests (attention on small N incl. stream-K partials, seeds, hypotheses, scoring, refinement, power iteration)"
timeout 1500 compute-sanitizer --tool $TOOL --error-exitcode 99 python -m pytest tests/test_gpu_kernels.py -x -q -m gpu \
  -k "(attention_split16_vs_oracle and (129 or 257) and not split16pv2) or pick_seeds_vs_reference or seed_knn_vs_reference or seed_hypotheses_vs_reference or score_hypotheses or post_refine or power_iter_full or mlp_chain or feature_head" 2>&1 | tail -12
echo "exit code ${PIPESTATUS[0]}"
} > $LOG 2>&1
tail -30 $LOG
