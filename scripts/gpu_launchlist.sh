#!/bin/bash
# ncu launch list (gpu__time_duration) of ONE steady-state forward -> gpurun_out/launches_forward.csv
mkdir -p gpurun_out
timeout 600 python scripts/prof_forward.py ${FWD_ARGS} > gpurun_out/prof_forward.log 2>&1 &&
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_forward.csv python scripts/prof_forward.py ${FWD_ARGS} > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches exit $?"
cat gpurun_out/prof_forward.log
