#!/bin/bash
# End-of-round visit on N GPUs (gpurun --gpus N): P2P power-iteration tests, bench.py under torchrun (incl. config 4), sweep cells.
N=${NGPU:-2}
mkdir -p gpurun_out
echo "== multi-GPU tests"; timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_forward.py -x -q -m gpu -k "p2p or two_devices or device" 2>&1 | tail -4
echo "== bench x$N"; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "exit $?"; cut -c1-300 gpurun_out/bench_n$N.json; python -c "
import json; d=json.load(open('gpurun_out/bench_n$N.json')); print('value', d['value'], 'e2e', d['e2e']['value'], 'config4', d.get('config4',{}).get('ms'), d.get('config4',{}).get('bit_identical_to_1gpu'), 'parity', d['parity_check']['ok'])"
NGPU=$N C4=${C4:-1} bash scripts/gpu_sweep.sh
