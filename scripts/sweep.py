"""BASELINE configs[4]: throughput sweep, 4096 registration pairs x N in {1000, 5000, 10000}, data-parallel over the GPUs
of one box, against the reference CPU path (pattern of the reference's test_parallel.sh:18-24 / test.py:236-257: every
process takes its share of the pairs, no communication).

One cell per invocation:   python scripts/sweep.py --num-corr 5000                       (1 GPU)
                           python -m torch.distributed.run --nproc-per-node 4 scripts/sweep.py --num-corr 5000
Pairs: seeds 1000 .. 5095, 3DMatch parameters, rank r takes the contiguous block parallel.shard_pairs gives it.
The inputs live in pinned HOST memory; the timed region is the whole job of the rank: per step H2D of one batch, forward,
D2H of final_trans + final_labels (CUDA events, max over ranks).  Rank 0 also runs the CPU port (oracle/restate.py, all host
threads) on the first pairs of the cell (16 / 8 / 4 pairs at N = 1000 / 5000 / 10000) and compares the
GPU results of those same pairs with it.  One JSON line on rank 0 (also written to gpurun_out/sweep/)."""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench
from oracle import restate, synth
from pointdsc_b200 import PointDSC, _cabi, parallel

ap = argparse.ArgumentParser()
ap.add_argument("--num-corr", dest="n", type=int, required=True)
ap.add_argument("--total-pairs", type=int, default=4096)
ap.add_argument("--batch", type=int, default=0, help="pairs per step per GPU (default: 64 / 16 / 8 at N = 1000 / 5000 / 10000)")
ap.add_argument("--cpu-pairs", type=int, default=-1)
args = ap.parse_args()
N = args.n
batch = args.batch or (64 if N <= 2000 else (16 if N <= 6000 else 8))
cpu_pairs = args.cpu_pairs if args.cpu_pairs >= 0 else (16 if N <= 2000 else (8 if N <= 6000 else 4))
world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
if world > 1 and args.cpu_pairs < 0:
    cpu_pairs = min(cpu_pairs, 2)          # the other GPUs idle while rank 0 runs the CPU port: keep the multi-GPU cells short
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist = None
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)
_cabi.check_device()
cfg = synth.CONFIGS["3dmatch"]
L = 12
sd = synth.make_state_dict(99, num_layers=L, sigma_d=cfg["sigma_d"])
model = PointDSC(num_layers=L, inlier_threshold=cfg["inlier_threshold"], sigma_d=cfg["sigma_d"], nms_radius=cfg["nms_radius"]).eval()
model.load_state_dict(sd, strict=True)
model = model.to(dev)

mine = parallel.shard_pairs(args.total_pairs, rank, world)
seeds = [1000 + i for i in mine]
keys = ("corr_pos", "src_keypts", "tgt_keypts")
t0 = time.perf_counter()
host = {k: torch.empty(len(seeds), N, 6 if k == "corr_pos" else 3).pin_memory() for k in keys}
for i, s in enumerate(seeds):
    p = synth.make_pair(s, N, **cfg)
    for k in keys:
        host[k][i] = p[k][0]
gen_s = time.perf_counter() - t0
out_trans = torch.empty(len(seeds), 4, 4).pin_memory()
out_labels = torch.empty(len(seeds), N).pin_memory()
steps = (len(seeds) + batch - 1) // batch

attn_events = []
orig = _cabi.nonlocal_attn_packed
def timed_attn(*a, **kw):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); out = orig(*a, **kw); e1.record()
    attn_events.append((e0, e1))
    return out
_cabi.nonlocal_attn_packed = timed_attn

def run_job():
    for s in range(steps):
        a, b = s * batch, min(len(seeds), (s + 1) * batch)
        data = {k: host[k][a:b].to(dev, non_blocking=True) for k in keys}
        data["testing"] = True
        r = model(data)
        out_trans[a:b].copy_(r["final_trans"], non_blocking=True)
        out_labels[a:b].copy_(r["final_labels"], non_blocking=True)
    torch.cuda.current_stream().synchronize()

def barrier():
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()

sampler = bench.ClockSampler(local)
sampler.prime(); sampler.start()
# warm-up: a slice of the job (same code path), until the clocks have settled
t_end = time.perf_counter() + 1.0
while time.perf_counter() < t_end:
    for s in range(min(steps, 4)):
        data = {k: host[k][s * batch:(s + 1) * batch].to(dev, non_blocking=True) for k in keys}
        data["testing"] = True
        model(data)
    torch.cuda.synchronize()
barrier()
attn_events.clear()
sampler.recording = True
l0 = _cabi.kernel_launches()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
run_job()
e1.record()
barrier()
launches = _cabi.kernel_launches() - l0
clocks = sampler.stop()
ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
if dist is not None:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
ms = float(ms[0])
attn_ms = [a.elapsed_time(b) for a, b in attn_events]

if rank == 0:
    pk = bench.peaks()
    # attention roofline: full batches only (the last step of a rank may be short)
    per_launch = sum(attn_ms) / max(1, len(attn_ms))
    flops = 4.0 * N * N * bench.C * batch
    achieved = flops / (per_launch * 1e-3) / 1e12
    line = {"metric": bench.METRIC, "value": args.total_pairs / (ms * 1e-3), "unit": bench.UNIT, "n_gpus": world,
            "config": {"workload": f"sweep cell: {args.total_pairs} pairs x N={N} (3DMatch-shaped, seeds 1000..{999 + args.total_pairs}), "
                                   f"{batch} pairs per step per GPU, 12 layers, seeded weights (seed 99)", "N": N, "pairs_total": args.total_pairs,
                       "pairs_per_step_per_gpu": batch, "steps_per_gpu": steps, "parallelism": f"pairs sharded over {world} GPU(s), no communication",
                       "inputs": "pinned host memory; H2D, forward and D2H of every step inside the timed region"},
            "job_ms": ms, "scaling": "strong (4096 pairs in total)", "gpu_launches": launches, "clocks": clocks,
            "h2d_bytes_per_step": batch * N * 12 * 4, "d2h_bytes_per_step": batch * (16 + N) * 4, "input_generation_s": gen_s,
            "roofline": {"bound": "tensor", "kernel": "pdsc_nonlocal_attn_packed[split16]", "achieved": achieved, "peak": pk["bf16_sustained"],
                         "unit": "TFLOP/s", "frac": achieved / pk["bf16_sustained"], "avg_launch_ms": per_launch, "launches_timed": len(attn_ms),
                         "share_of_job": sum(attn_ms) / ms, "peak_source": pk["source"] + ", bf16 sustained", "mma_terms_avg": 3.0,
                         "issued_frac": 3.0 * achieved / pk["bf16_sustained"]}}
    if cpu_pairs > 0:
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        times, worst = [], dict(dconf=0.0, rot=0.0, trans=0.0, label_mismatch=0)
        for i in range(cpu_pairs + 1):                     # first one is the warm-up
            j = min(i, len(seeds) - 1)
            t0 = time.perf_counter()
            ref = restate.forward_pair(sd, host["corr_pos"][j:j + 1], host["src_keypts"][j:j + 1], host["tgt_keypts"][j:j + 1], num_layers=L,
                                       inlier_threshold=cfg["inlier_threshold"], nms_radius=cfg["nms_radius"])
            if i > 0:
                times.append(time.perf_counter() - t0)
            # parity of the GPU job's result for this pair (confidences are not kept by the job; transform and labels are)
            re = float(restate.rotation_error_deg(out_trans[j:j + 1, :3, :3], ref["final_trans"][:, :3, :3]))
            te = float(restate.translation_error(out_trans[j:j + 1, :3, 3], ref["final_trans"][:, :3, 3]))
            worst["rot"], worst["trans"] = max(worst["rot"], re), max(worst["trans"], te)
            worst["label_mismatch"] = max(worst["label_mismatch"], int((out_labels[j] != ref["final_labels"][0]).sum()))
        mean = sum(times) / len(times)
        line["cpu_baseline"] = {"value": 1.0 / mean, "unit": bench.UNIT, "cores": cores, "kind": "port", "s_per_pair": mean,
                                "sample": f"{cpu_pairs} pair(s) of this cell after 1 warm-up pair, oracle/restate.py, torch CPU fp32, {cores} threads"}
        if world > 1:
            line["cpu_baseline"]["note"] = ("timed while the other ranks' host threads spin in the closing barrier: NOT a baseline; "
                                            "the CPU figure of a row of the sweep is the 1-GPU cell's")
        line["parity_check"] = {"against": f"oracle/restate.py on the first {cpu_pairs + 1} pairs of the cell", "max_rot_err_deg": worst["rot"],
                                "max_trans_err_m": worst["trans"], "max_label_mismatches": worst["label_mismatch"],
                                "ok": worst["rot"] <= 0.01 and worst["trans"] <= 1e-4 and worst["label_mismatch"] <= 2,
                                "tolerance": "rot <= 0.01 deg, trans <= 1e-4 m; labels exact apart from exact-threshold ties (<= 2)"}
    os.makedirs(os.path.join(ROOT, "gpurun_out", "sweep"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "sweep", f"cell_n{N}_g{world}.json"), "w") as f:
        json.dump(line, f, indent=1)
    print(json.dumps(line), flush=True)
if dist is not None:
    dist.barrier()
    dist.destroy_process_group()
