#!/bin/bash
# which MMAs cost what: attention launch time with subsets of the hi/lo terms (tuning library; results are wrong by design)
run() { echo "== $*"; env "$@" python scripts/prof_attn.py split16 16 5000 20 | tail -1; }
python scripts/prof_attn.py split16 16 5000 20 | tail -1
run PDSC_ATTN_GRID=148
run PDSC_ATTN_QK_TERMS=3
run PDSC_ATTN_QK_TERMS=5
run PDSC_ATTN_QK_TERMS=1
run PDSC_ATTN_PV_TERMS=2
run PDSC_ATTN_PV_TERMS=1
run PDSC_ATTN_QK_TERMS=1 PDSC_ATTN_PV_TERMS=1
run PDSC_ATTN_DBG=1
run PDSC_ATTN_DBG=2
run PDSC_ATTN_DBG=3
run PDSC_ATTN_DBG=4
run PDSC_ATTN_DBG=8
