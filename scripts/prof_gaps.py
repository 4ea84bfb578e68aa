"""Kernel timeline of one steady-state forward (torch.profiler / CUPTI activity records): busy time, idle gaps between
consecutive kernels on the stream, and the largest gaps.  python scripts/prof_gaps.py [pairs] [N] [family]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
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
for _ in range(5):
    model(data)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        model(data)
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.time_range is not None]
ks = sorted(((e.time_range.start, e.time_range.end, e.name) for e in ev), key=lambda t: t[0])
n = len(ks) // 3
ks = ks[-n:]                                   # the last forward
busy = sum(e - s for s, e, _ in ks)
span = ks[-1][1] - ks[0][0]
gaps = [(ks[i + 1][0] - ks[i][1], ks[i][2][:40], ks[i + 1][2][:40]) for i in range(len(ks) - 1)]
out = {"launches": len(ks), "span_us": span, "busy_us": busy, "idle_us": span - busy,
       "median_gap_us": sorted(g[0] for g in gaps)[len(gaps) // 2],
       "largest_gaps": [{"us": g[0], "after": g[1], "before": g[2]} for g in sorted(gaps, reverse=True)[:8]]}
print(json.dumps(out, indent=1))
