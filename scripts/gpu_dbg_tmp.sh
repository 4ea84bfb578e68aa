for d in 8 24; do echo "== PDSC_ATTN_DBG=$d"; for n in 256 384 512; do PDSC_ATTN_DBG=$d timeout 100 python scripts/dbg_attn.py $n 0,0,0,3,7 2>&1 | tail -5 | cut -c1-75; done; done
