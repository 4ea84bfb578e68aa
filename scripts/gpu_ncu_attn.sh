#!/bin/bash
# ncu --set full capture of the attention main kernel alone (scripts/prof_attn.py), report -> gpurun_out/
mkdir -p gpurun_out
K=${NCU_KERNEL:-attn_pp_kernel}
TAG=${NCU_TAG:-attn_pp}
timeout 300 python scripts/prof_attn.py ${PROF_ARGS:-split16 16 5000 3} > gpurun_out/prof_attn.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 2 -c 1 -f -o gpurun_out/prof_$TAG python scripts/prof_attn.py ${PROF_ARGS:-split16 16 5000 3} > gpurun_out/ncu_full.log 2>&1; echo "ncu full exit $?"
cat gpurun_out/prof_attn.log; tail -3 gpurun_out/ncu_full.log
