"""GraphedForward against the eager forward: results bit for bit, time per forward.  python scripts/prof_graph.py [pairs] [N]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import synth
from pointdsc_b200 import PointDSC
from pointdsc_b200.graphed import GraphedForward

pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 64
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
cfg = synth.CONFIGS["3dmatch"]
sd = synth.make_state_dict(99, num_layers=12, sigma_d=cfg["sigma_d"])
model = PointDSC(num_layers=12, inlier_threshold=cfg["inlier_threshold"], sigma_d=cfg["sigma_d"], nms_radius=cfg["nms_radius"]).eval()
model.load_state_dict(sd, strict=True)
model = model.cuda()
batches = [synth.make_batch(list(range(s, s + pairs)), N, **cfg) for s in (0, 500)]
dev = [{k: b[k].cuda() for k in ("corr_pos", "src_keypts", "tgt_keypts")} for b in batches]
fwd = GraphedForward(model, pairs, N)
ok = True
for d in dev + dev[:1]:
    eager = model(dict(d, testing=True))
    g = fwd(d["corr_pos"], d["src_keypts"], d["tgt_keypts"])
    torch.cuda.synchronize()
    for k in ("confidence", "seeds", "final_trans", "final_labels", "inlier_count"):
        ok &= bool(torch.equal(eager[k], g[k]))
    ok &= bool(g["finite"])
print("graph replay bit-identical to the eager forward on 3 batches:", ok)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
def timeit(fn, reps=30):
    for _ in range(5): fn()
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
d = dev[0]
for rnd in range(3):
    te = timeit(lambda: model(dict(d, testing=True)))
    tg = timeit(lambda: fwd(d["corr_pos"], d["src_keypts"], d["tgt_keypts"]))
    print(f"{pairs} x N={N}: eager {te:.3f} ms, graph {tg:.3f} ms per forward ({100 * (te - tg) / te:.1f} % less)")
