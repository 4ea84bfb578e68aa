bash scripts/gpu_launchlist.sh
timeout 600 python bench.py --family kitti --n 10000 --no-config4 --no-cpu-baseline > gpurun_out/bench_final_kitti.json 2> gpurun_out/bench_final_kitti.err; echo "exit $?"; python -c "
import json; d=json.load(open('gpurun_out/bench_final_kitti.json')); print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['avg_launch_ms'], d['parity_check']['ok'], d['parity_check'].get('ill_conditioned_pairs'))"
