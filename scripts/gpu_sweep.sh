#!/bin/bash
# config 5 sweep for the GPU count of this gpurun call: NGPU=1|2|4|8 bash scripts/gpu_sweep.sh   (+ config 4 at that GPU count)
G=${NGPU:-1}
mkdir -p gpurun_out/sweep
for n in 1000 5000 10000; do
  if [ "$G" = "1" ]; then timeout 900 python scripts/sweep.py --num-corr $n 2> gpurun_out/sweep/cell_n${n}_g$G.err | cut -c1-400
  else timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 295$((10 + n / 1000)) scripts/sweep.py --num-corr $n 2> gpurun_out/sweep/cell_n${n}_g$G.err | cut -c1-400; fi
  echo "cell n=$n g=$G exit ${PIPESTATUS[0]}"
done
if [ "${C4:-1}" = "1" ] && [ "$G" != "1" ]; then
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29540 scripts/shard_power.py 50000 > gpurun_out/shard_power_n$G.json 2> gpurun_out/shard_power_n$G.err; echo "shard_power x$G exit $?"; grep "^{" gpurun_out/shard_power_n$G.json | cut -c1-900
else
  timeout 300 python scripts/shard_power.py 50000 | grep "^{" > gpurun_out/shard_power_n1.json; cat gpurun_out/shard_power_n1.json | cut -c1-500
fi
