#!/bin/bash
mkdir -p gpurun_out
echo "== no-alloc test alone, fresh process"; timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -3
echo "== pair 10"; timeout 300 python scripts/dump_seed.py batch_3dmatch_n5000_l12 10 2>&1 | tail -22
echo "== whole gpu suite"; timeout 2400 python -m pytest tests -x -q -m gpu 2>&1 | tail -15
cp gpurun_out/parity_report_*.json gpurun_out/ 2>/dev/null; ls gpurun_out
echo "== kitti budget"; timeout 900 python scripts/kitti_budget.py 2>&1 | tail -14
