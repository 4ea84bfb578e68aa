timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "exit $?"; tail -2 gpurun_out/bench_n2.err
python -c "
import json; d=json.load(open('gpurun_out/bench_n2.json')); print(d['value'], d['step_ms']); print(d['e2e']['value'], d['e2e']['step_ms']); print(d['parity_check']['ok'], d['gpu_launches'], d['config']['untimed_rehearsal_regions'], d['config4']['ms'], d['config4']['bit_identical_to_1gpu'])"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 | cut -c1-200
