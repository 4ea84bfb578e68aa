#!/bin/bash
# Front-end round: GPU parity tests of N1, then the stage timings.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_frontend.py -m gpu -q -rA --timeout=600 > gpurun_out/pytest_frontend.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/pytest_frontend.log
grep -v "^PASSED" gpurun_out/pytest_frontend.log | tail -60
timeout 600 python scripts/bench_frontend.py > gpurun_out/frontend.json 2> gpurun_out/frontend.err; echo "bench exit $?"
cat gpurun_out/frontend.json; tail -5 gpurun_out/frontend.err
