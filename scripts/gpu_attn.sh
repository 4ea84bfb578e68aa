#!/bin/bash
# Attention-kernel iteration on the GPU box: its parity tests, then timing of the kernel alone.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "attention" --timeout=300 > gpurun_out/pytest_attn.log 2>&1
echo "pytest exit $?"; tail -15 gpurun_out/pytest_attn.log
for g in ${GRIDS:-sk 0}; do
  if [ "$g" = "sk" ]; then unset PDSC_ATTN_GRID; else export PDSC_ATTN_GRID=$g; fi
  echo "== GRID=$g"
  timeout 120 python scripts/prof_attn.py split16 16 5000 10
  timeout 120 python scripts/prof_attn.py fp16 16 5000 10
done
unset PDSC_ATTN_GRID

