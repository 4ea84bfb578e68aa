#!/bin/bash
# End-of-round visit on one GPU: whole GPU suite, smoke, both bench arms, the 1-GPU cells of the sweep.
mkdir -p gpurun_out
echo "== gpu suite"; timeout 2400 python -m pytest tests -x -q -m gpu 2>&1 | tail -6
echo "== smoke"; timeout 600 python __graft_entry__.py --smoke 2>&1 | tail -3
echo "== bench"; timeout 900 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "exit $?"; cut -c1-700 gpurun_out/bench_final.json
echo "== bench reference"; timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_final_reference.json 2>> gpurun_out/bench_final.err; echo "exit $?"; cut -c1-300 gpurun_out/bench_final_reference.json
echo "== kitti bench"; timeout 900 python bench.py --family kitti --n 10000 --no-config4 --no-cpu-baseline > gpurun_out/bench_final_kitti.json 2>> gpurun_out/bench_final.err; echo "exit $?"; cut -c1-400 gpurun_out/bench_final_kitti.json
NGPU=1 C4=0 bash scripts/gpu_sweep.sh
