#!/bin/bash
# Full GPU test suite + smoke + the default bench line + the other BASELINE configs as extra bench lines.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout=900 > gpurun_out/pytest_all.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_all.log
tail -3 gpurun_out/pytest_all.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
timeout 900 python bench.py --n 1000 --pairs 64 --no-cpu-baseline > gpurun_out/bench_n1000.json 2>> gpurun_out/bench.err; echo "bench n1000 exit $?"
timeout 900 python bench.py --n 10000 --pairs 16 --family kitti --cpu-pairs 2 > gpurun_out/bench_kitti_n10000.json 2>> gpurun_out/bench.err; echo "bench kitti exit $?"
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2>> gpurun_out/bench.err; echo "bench ref exit $?"
for f in bench bench_n1000 bench_kitti_n10000 bench_reference; do python - <<PY
import json
d=json.load(open("gpurun_out/$f.json"))
print("$f", d.get("value"), d.get("unit"), "e2e", d.get("e2e",{}).get("value"), "roof", d.get("roofline",{}).get("frac"), "cpu", d.get("cpu_baseline",{}).get("value"))
PY
done
tail -3 gpurun_out/bench.err
