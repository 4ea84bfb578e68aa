#!/bin/bash
# round 2, first GPU check: attention tests on the new term set, per-launch timing of both term sets, a short bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "attention or knn" 2>&1 | tail -5 > gpurun_out/s1_tests.log
for m in split16 split16x3; do timeout 120 python scripts/prof_attn.py $m 16 5000 20; done > gpurun_out/s1_prof.log 2>&1
timeout 120 python scripts/prof_attn.py split16 64 1000 20 >> gpurun_out/s1_prof.log 2>&1
timeout 120 python scripts/prof_attn.py split16 16 10000 10 >> gpurun_out/s1_prof.log 2>&1
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/s1_bench.json 2> gpurun_out/s1_bench.err
cat gpurun_out/s1_tests.log gpurun_out/s1_prof.log; head -c 1500 gpurun_out/s1_bench.json
