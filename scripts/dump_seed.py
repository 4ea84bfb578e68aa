"""Why does one pair's best-seed hypothesis differ?  Runs the oracle for ONE pair of a batch golden and compares the
intermediates of the winning seed: python scripts/dump_seed.py batch_3dmatch_n5000_l12 10   (GPU + ~1 s of CPU oracle)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from conftest import load_golden
from oracle import restate, synth
from parity import golden_model
from pointdsc_b200 import _cabi
case, i = sys.argv[1], int(sys.argv[2])
meta, g = load_golden(case)
fam = synth.CONFIGS[meta["family"]]
model, sd = golden_model(meta)
p = synth.make_pair(meta["pair_seeds"][i], meta["N"], **fam)
ref = restate.forward_pair(sd, p["corr_pos"], p["src_keypts"], p["tgt_keypts"], num_layers=meta["num_layers"],
                           inlier_threshold=fam["inlier_threshold"], nms_radius=fam["nms_radius"])
data = {k: p[k].cuda() for k in ("corr_pos", "src_keypts", "tgt_keypts")}; data["testing"] = True
res = model(data)
b = int(ref["best_idx"][0]) if "best_idx" in ref else int(torch.argmax(ref["seed_fitness"][0]))
seed = int(ref["seeds"][0, b])
ob = int(res["best_idx"][0]); print("ref best slot", b, "seed", seed, "| ours best slot", ob, "seed", int(res["seeds"][0, ob]))
ok = ref["knn_idx"][0, b]; mine = res["knn_idx"][0, ob].cpu()
print("knn lists equal:", bool(torch.equal(ok, mine)), "| same sets:", set(ok.tolist()) == set(mine.tolist()))
if not torch.equal(ok, mine):
    print(" ref :", ok.tolist()); print(" ours:", mine.tolist())
    f = ref["normed_features"][0]
    d = 2 - 2 * (f[seed][None] @ f.T)[0]
    for j in sorted(set(ok.tolist()) ^ set(mine.tolist())): print("   idx", j, "fp32 ref dist", float(d[j]))
    print("   40th/41st smallest ref distances:", torch.sort(d)[0][39:43].tolist())
# run the hypothesis kernel on the REFERENCE's features / knn list for this seed and compare
nf = ref["normed_features"].contiguous().cuda(); kn = ref["knn_idx"].contiguous().cuda()
tr, w, it = _cabi.seed_hypotheses(nf, data["src_keypts"], data["tgt_keypts"], kn, 1.0, fam["sigma_d"], 10, want_aux=True)
rw = ref["weights"].reshape(w.shape[1], -1)
print("on reference inputs: weights max diff (best seed)", float((w[0, b].cpu() - rw[b]).abs().max()), "trans diff", float((tr[0, b].cpu() - ref["seed_trans"][0, b]).abs().max()))
print("our pipeline: seed_trans diff", float((res["seed_trans"][0, ob].cpu() - ref["seed_trans"][0, b]).abs().max()))
A = p["src_keypts"][0][ok].double(); Bp = p["tgt_keypts"][0][ok].double(); wr = rw[b].double()
wn = wr / (wr.sum() + 1e-6); cA = (wn[:, None] * A).sum(0); cB = (wn[:, None] * Bp).sum(0)
H = (A - cA).T @ torch.diag(wr) @ (Bp - cB); print("singular values of H:", torch.linalg.svdvals(H).tolist(), "weights: max", float(wr.max()), "n>1e-3", int((wr > 1e-3).sum()))
print("normed feature diff at the knn rows:", float((res["corr_features"][0].cpu()[ok] / res["corr_features"][0].cpu()[ok].norm(dim=-1, keepdim=True) - ref["normed_features"][0][ok]).abs().max()))
