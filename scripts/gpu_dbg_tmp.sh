run() { echo "== $*"; env "$@" python scripts/prof_attn.py split16 16 5000 20 | tail -1 | cut -c1-90; }
run PDSC_ATTN_DBG=8
run PDSC_ATTN_DBG=1
run PDSC_ATTN_DBG=2
run PDSC_ATTN_DBG=4
run PDSC_ATTN_DBG=5
run PDSC_ATTN_QK_TERMS=1
run PDSC_ATTN_QK_TERMS=3
run PDSC_ATTN_QK_TERMS=5
run PDSC_ATTN_PV_TERMS=1
