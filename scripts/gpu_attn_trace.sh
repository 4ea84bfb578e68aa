#!/bin/bash
# Timelines of CTA 0 (debug instantiation of the attention kernel) under the ablation knobs.
mkdir -p gpurun_out
for dbg in 8 7 4 3 1; do
  PDSC_ATTN_DBG=$dbg PDSC_ATTN_TRACE=gpurun_out/attn_trace_dbg$dbg.txt timeout 120 python scripts/prof_attn.py split16 16 5000 1
done
