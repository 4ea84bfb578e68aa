#!/bin/bash
# Multi-GPU visit (gpurun --gpus N): row-sharded power iteration (config 4, P2P + all-gather) and the pair-sharded bench.
N=${NGPU:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
nvidia-smi topo -m 2>/dev/null | head -12
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 scripts/shard_power.py 50000 > gpurun_out/shard_power_n$N.json 2> gpurun_out/shard_power_n$N.err; echo "shard_power x$N exit $?"
tail -2 gpurun_out/shard_power_n$N.json; tail -5 gpurun_out/shard_power_n$N.err
if [ "${BENCH:-1}" = "1" ]; then
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench x$N exit $?"
cat gpurun_out/bench_n$N.json | cut -c1-3000; tail -3 gpurun_out/bench_n$N.err
fi
