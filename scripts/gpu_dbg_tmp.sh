python scripts/ab_attn.py 16 5000 20 head.so final.so qcoal.so
python scripts/ab_attn.py 64 1000 40 head.so final.so qcoal.so
python scripts/ab_attn.py 16 10000 8 head.so final.so qcoal.so
timeout 900 python -m pytest tests/test_gpu_kernels.py -x -q -k "attn or attention" 2>&1 | tail -3
