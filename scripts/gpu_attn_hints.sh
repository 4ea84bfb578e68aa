#!/bin/bash
run() { echo "== $*"; env "$@" python scripts/prof_attn.py split16 16 5000 20 | tail -1 | cut -c1-90; }
python scripts/prof_attn.py split16 16 5000 20 | tail -1 | cut -c1-90
for h in 0 100 1000; do run PDSC_ATTN_HINT_SOFT=$h; run PDSC_ATTN_HINT_ROLE=$h; run PDSC_ATTN_HINT_SOFT=$h PDSC_ATTN_HINT_ROLE=$h; done
run PDSC_ATTN_HINT_SOFT=0 PDSC_ATTN_HINT_ROLE=0 PDSC_ATTN_DBG=4
run PDSC_ATTN_HINT_SOFT=20000 PDSC_ATTN_DBG=4
