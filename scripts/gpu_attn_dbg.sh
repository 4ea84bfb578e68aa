#!/bin/bash
for dbg in 0 1 2 3 4 5 6 7; do
  echo "== DBG=$dbg (1=no compat, 2=no exp, 4=no MMA)"
  PDSC_ATTN_DBG=$dbg timeout 120 python scripts/prof_attn.py split16 16 5000 10
done
