"""Times the ICP hand-off on the GPU and the CPU restatement on a bounded sample.
python scripts/bench_icp.py > profiles/r01_icp.json"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from pointdsc_b200 import _cabi
from oracle import restate_icp as I, synth

B, N = 16, 5000
cfg = synth.CONFIGS["3dmatch"]
batch = synth.make_batch(list(range(B)), N, **cfg)
src, tgt = batch["src_keypts"].cuda(), batch["tgt_keypts"].cuda()
T0 = batch["gt_trans"].clone()
T0[:, :3, 3] += 0.02
T0 = T0.cuda()
for _ in range(3):
    _cabi.icp_refine(src, tgt, T0)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
reps = 10
for _ in range(reps):
    out, stats, iters, _ = _cabi.icp_refine(src, tgt, T0)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
t0 = time.perf_counter()
for b in range(2):
    I.icp_point_to_point(src[b].cpu().numpy(), tgt[b].cpu().numpy(), T0[b].cpu().numpy())
cpu = (time.perf_counter() - t0) / 2
print(json.dumps({"workload": f"{B} pairs x {N} key points (synthetic 3dmatch-shaped), max distance 0.1, <= 30 iterations",
                  "gpu_ms_per_call": ms, "pairs_per_s": B / ms * 1e3, "updates_applied": iters.tolist(),
                  "launches_per_call": 1 + 2 * 31, "cpu_port_s_per_pair": cpu, "cpu_kind": "port (oracle/restate_icp.py, scipy cKDTree)",
                  "speedup_vs_cpu_port": cpu / (ms / 1e3 / B)}, indent=1))
