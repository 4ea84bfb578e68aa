"""BASELINE config 4 alone: one pair with N correspondences, full-matrix power iteration with the compatibility rows
sharded over the ranks (bench.py's config4 leg: NVLink P2P stores + device flags, NCCL all-gather beside it).  Launch with
torchrun; prints one JSON line on rank 0.   python -m torch.distributed.run --nproc-per-node 2 scripts/shard_power.py 50000"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
import bench

N = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
leg = bench.config4_leg(N, dist if world > 1 else None, dev, rank, world)
if rank == 0:
    print(json.dumps(leg), flush=True)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
