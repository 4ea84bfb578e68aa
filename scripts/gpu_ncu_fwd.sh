#!/bin/bash
# ncu --set full of ONE launch of a named kernel inside a steady-state forward -> gpurun_out/prof_<tag>.ncu-rep
mkdir -p gpurun_out
K=${NCU_KERNEL:?kernel regex}; TAG=${NCU_TAG:-fwd}; SKIP=${NCU_SKIP:-2}
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:$K -s $SKIP -c 1 -f -o gpurun_out/prof_$TAG python scripts/prof_forward.py ${FWD_ARGS} > gpurun_out/ncu_fwd.log 2>&1; echo "ncu exit $?"; tail -2 gpurun_out/ncu_fwd.log
