#!/bin/bash
# Next-row round (N1 front end, N4 ICP): GPU parity tests, stage timings, ncu launch list of one front-end pass.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_frontend.py tests/test_gpu_icp.py -m gpu -q -rA --timeout=600 > gpurun_out/pytest_frontend.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/pytest_frontend.log
grep -v "^PASSED" gpurun_out/pytest_frontend.log | tail -${TAIL:-60}
timeout 600 python scripts/bench_frontend.py > gpurun_out/frontend.json 2> gpurun_out/frontend.err; echo "bench exit $?"
tail -5 gpurun_out/frontend.err
timeout 600 python scripts/bench_icp.py > gpurun_out/icp.json 2> gpurun_out/icp.err; echo "icp bench exit $?"
cat gpurun_out/icp.json; tail -5 gpurun_out/icp.err
if [ -n "${NCU}" ]; then
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'grid_|hybrid|voxel|spfh|fpfh|match|icp_' -c 400 --csv --log-file gpurun_out/launches_frontend.csv python scripts/bench_frontend.py > gpurun_out/ncu_frontend.log 2>&1; echo "ncu exit $?"
fi
