#!/bin/bash
# A/B timing of the attention kernel: production vs debug instantiation (dbg=8 sets no knob), both launch shapes.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "attention" --timeout=300 > gpurun_out/pytest_attn.log 2>&1
echo "pytest exit $?"; tail -3 gpurun_out/pytest_attn.log
for rep in 1 2; do
echo "== production"; timeout 120 python scripts/prof_attn.py split16 16 5000 20
echo "== production, cluster 2"; PDSC_ATTN_CLUSTER=2 timeout 120 python scripts/prof_attn.py split16 16 5000 20
echo "== debug instantiation"; PDSC_ATTN_DBG=8 timeout 120 python scripts/prof_attn.py split16 16 5000 20
done
