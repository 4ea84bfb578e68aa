#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench, ncu launch list of the bench, ncu full capture of the attention kernel.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.used --format=csv > gpurun_out/smi.txt 2>&1
if [ -z "${SKIP_TESTS}" ]; then
timeout 1500 python -m pytest tests -m gpu -q -rA --timeout=900 ${PYTEST_ARGS} > gpurun_out/pytest.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/pytest.log
tail -25 gpurun_out/pytest.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
fi
timeout 900 python bench.py ${BENCH_ARGS} > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
if [ -n "${NCU_LAUNCHES}" ]; then
timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c ${NCU_LAUNCHES} --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches exit $?"
fi
if [ -n "${NCU_FULL}" ]; then
timeout 300 python scripts/prof_attn.py ${PROF_ARGS} > gpurun_out/prof_attn.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:${NCU_FULL} -s 2 -c 1 -f -o gpurun_out/prof_${NCU_TAG:-attn} python scripts/prof_attn.py ${PROF_ARGS} > gpurun_out/ncu_full.log 2>&1; echo "ncu full exit $?"
cat gpurun_out/prof_attn.log
fi
