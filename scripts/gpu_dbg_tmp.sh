timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_forward.py -x -q -m gpu -k "dups or duplicated" -s 2>&1 | grep -v "^$" | tail -25
