"""Same-process A/B of the forward with and without the joined layer chains (model.join_layers), optionally for several
library variants (scripts/ab_variant.py --file mlp.cu ...): interleaved groups of forwards, CUDA-event time per forward,
results compared bit for bit.   python scripts/ab_forward.py [pairs] [N] [reps] [libA.so libB.so ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import synth
from pointdsc_b200 import PointDSC, _cabi

pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 16
N = int(sys.argv[2]) if len(sys.argv) > 2 else 5000
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 10
cfg = synth.CONFIGS["3dmatch"]
batch = synth.make_batch(list(range(pairs)), N, **cfg)
sd = synth.make_state_dict(99, num_layers=12, sigma_d=cfg["sigma_d"])
model = PointDSC(num_layers=12, inlier_threshold=cfg["inlier_threshold"], sigma_d=cfg["sigma_d"], nms_radius=cfg["nms_radius"]).eval()
model.load_state_dict(sd, strict=True)
model = model.cuda()
data = {k: batch[k].cuda() for k in ("corr_pos", "src_keypts", "tgt_keypts")}
data["testing"] = True
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
libs = [os.path.join(ROOT, "pointdsc_b200", "lib", "ab", x) for x in sys.argv[4:]]
configs = [(_cabi.LIB_PATH, False), (_cabi.LIB_PATH, True)] + [(l, True) for l in libs]
times, outs = {c: [] for c in configs}, {}
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for rnd in range(7):
    for cfg_ in configs:
        _cabi._active, join = cfg_
        model.join_layers = join
        for _ in range(3 if rnd == 0 else 1):
            r = model(data)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            r = model(data)
        e1.record()
        torch.cuda.synchronize()
        if rnd:
            times[cfg_].append(e0.elapsed_time(e1) / reps)
        outs[cfg_] = r
for cfg_ in configs:
    t = sorted(times[cfg_])
    same = all(torch.equal(outs[cfg_][k], outs[configs[0]][k]) for k in ("final_trans", "final_labels", "confidence", "seeds"))
    print(f"{pairs} x N={N} {os.path.basename(cfg_[0]):26s} join_layers={cfg_[1]!s:5s}: median {t[len(t) // 2]:.3f} ms per forward "
          f"(min {t[0]:.3f}, max {t[-1]:.3f}); identical to the first: {same}")
