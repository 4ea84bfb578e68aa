"""One steady-state forward of the bench workload between cudaProfilerStart/Stop (for `ncu --profile-from-start off`).
python scripts/prof_forward.py [pairs] [N] [family]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import synth
from pointdsc_b200 import PointDSC

pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 16
N = int(sys.argv[2]) if len(sys.argv) > 2 else 5000
family = sys.argv[3] if len(sys.argv) > 3 else "3dmatch"
cfg = synth.CONFIGS[family]
batch = synth.make_batch(list(range(pairs)), N, **cfg)
sd = synth.make_state_dict(99, num_layers=12, sigma_d=cfg["sigma_d"])
model = PointDSC(num_layers=12, inlier_threshold=cfg["inlier_threshold"], sigma_d=cfg["sigma_d"], nms_radius=cfg["nms_radius"]).eval()
model.load_state_dict(sd, strict=True)
model = model.cuda()
data = {k: batch[k].cuda() for k in ("corr_pos", "src_keypts", "tgt_keypts")}
data["testing"] = True
for _ in range(3):
    model(data)
torch.cuda.synchronize()
t0 = time.perf_counter()
torch.cuda.cudart().cudaProfilerStart()
res = model(data)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print(f"forward of {pairs} pairs x N={N}: {1e3 * (time.perf_counter() - t0):.2f} ms wall")
