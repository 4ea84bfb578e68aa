for i in 1 2; do
timeout 900 python bench.py --no-config4 --no-cpu-baseline > gpurun_out/bench_e2e_test$i.json 2> gpurun_out/bench_e2e_test$i.err; echo "exit $?"; tail -2 gpurun_out/bench_e2e_test$i.err
python -c "
import json; d=json.load(open('gpurun_out/bench_e2e_test$i.json')); print(d['value'], d['step_ms']); print(d['e2e']['value'], d['e2e']['step_ms']); print(d['parity_check']['ok'], d['gpu_launches'], d['config']['untimed_rehearsal_regions'], d['config']['extra_warmup_steps_until_steady'], d['roofline']['frac'])"
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 1 --no-config4 --no-cpu-baseline | cut -c1-120
