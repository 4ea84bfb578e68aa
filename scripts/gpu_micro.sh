#!/bin/bash
# micro-benchmarks of the CUDA-core side of the attention kernel (results -> gpurun_out/micro.log)
mkdir -p gpurun_out
for f in mufu_rate softmax_rate lds_rate; do echo "== $f"; timeout 60 scripts/micro/$f; done 2>&1 | tee gpurun_out/micro.log
