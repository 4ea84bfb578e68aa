#!/bin/bash
# Cluster-of-2 (TMA multicast) attention variant: parity tests, then timing against the plain launch.
mkdir -p gpurun_out
export PDSC_ATTN_CLUSTER=2
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "attention" --timeout=300 > gpurun_out/pytest_attn_cl2.log 2>&1
echo "pytest(cluster=2) exit $?"; tail -15 gpurun_out/pytest_attn_cl2.log
for cl in 2 1; do
  export PDSC_ATTN_CLUSTER=$cl
  echo "== CLUSTER=$cl"
  timeout 120 python scripts/prof_attn.py split16 16 5000 10
  timeout 120 python scripts/prof_attn.py fp16 16 5000 10
  PDSC_ATTN_DBG=7 timeout 120 python scripts/prof_attn.py split16 16 5000 10
done
