#!/bin/bash
# Runs on the GPU box under gpurun: tests, smoke, bench, ncu launch list.  Outputs -> gpurun_out/
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.used --format=csv > gpurun_out/smi.txt 2>&1
echo "=== pytest -m gpu" 
timeout 1200 python -m pytest tests -m gpu -q -rA --timeout=900 ${PYTEST_ARGS} > gpurun_out/pytest.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/pytest.log
tail -60 gpurun_out/pytest.log
echo "=== smoke"
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/smoke.log
tail -5 gpurun_out/smoke.log
echo "=== bench"
timeout 900 python bench.py ${BENCH_ARGS} > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
if [ -n "${NCU_LAUNCHES}" ]; then
echo "=== ncu launch list"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c ${NCU_LAUNCHES} --csv --log-file gpurun_out/launches.csv python bench.py ${NCU_BENCH_ARGS} > gpurun_out/ncu.log 2>&1; echo "ncu exit $?"
tail -3 gpurun_out/ncu.log
fi
