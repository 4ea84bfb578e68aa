"""BASELINE config 4: one pair with N=50000 correspondences, full-matrix power iteration with the compatibility
rows sharded over the ranks (NCCL all-gather of the eigenvector per iteration).  Launch with torchrun.
Checks the sharded result against the single-GPU driver on rank 0 and prints timings as one JSON line."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from oracle import synth
from pointdsc_b200 import _cabi, parallel

N = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
cfg = synth.CONFIGS["kitti"]
pair = synth.make_pair(200, N, **cfg)
src, tgt = pair["src_keypts"][0].to(dev), pair["tgt_keypts"][0].to(dev)
for _ in range(2):
    v, iters = parallel.power_iteration_full_sharded(src, tgt, cfg["sigma_d"], 10)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
v, iters = parallel.power_iteration_full_sharded(src, tgt, cfg["sigma_d"], 10)
e1.record()
torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    v1, it1 = _cabi.power_iter_full(src, tgt, cfg["sigma_d"], 10)
    torch.cuda.synchronize()
    err = float((v - v1).abs().max())
    n_it = int(iters.item())
    print(json.dumps({"config": f"single pair N={N}, row-sharded power iteration", "n_gpus": world, "ms": float(ms[0]), "iterations": n_it,
                      "iterations_single_gpu": int(it1.item()) if hasattr(it1, "item") else int(it1),
                      "max_abs_diff_vs_single_gpu": err, "effective_GBps_4N2_per_iter": 4.0 * N * N * n_it / (float(ms[0]) * 1e-3) / 1e9}), flush=True)
    assert err < 1e-6
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
