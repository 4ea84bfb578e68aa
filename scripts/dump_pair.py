"""Per-pair detail of the production batch against the 16-pair golden: python scripts/dump_pair.py [case] (GPU)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from conftest import load_golden
from oracle import synth
from parity import batch_parity_report, golden_model, point_distances
case = sys.argv[1] if len(sys.argv) > 1 else "batch_3dmatch_n5000_l12"
meta, g = load_golden(case)
fam = synth.CONFIGS[meta["family"]]
model, _ = golden_model(meta)
batch = synth.make_batch(meta["pair_seeds"], meta["N"], **fam)
data = {k: batch[k].cuda() for k in ("corr_pos", "src_keypts", "tgt_keypts")}
data["testing"] = True
res = model(data)
rep = batch_parity_report(case, meta, g, batch, res, fam)
thr = fam["inlier_threshold"]
for i, row in enumerate(rep["pairs"]):
    ours_best = int(res["seeds"][i][int(res["best_idx"][i])]); ref_best = int(g["seeds"][i][int(g["best_idx"][i])])
    oc = int(res["inlier_count"][i][int(res["best_idx"][i])]); rc = int(g["inlier_count"][i][int(g["best_idx"][i])])
    # the reference's count for OUR best seed (if the reference holds that seed)
    pos = (g["seeds"][i] == ours_best).nonzero().flatten()
    ref_count_of_ours = int(g["inlier_count"][i][pos[0]]) if pos.numel() else None
    n_at_max = int((g["inlier_count"][i] == rc).sum())
    from parity import trans_errors
    ire, ite = trans_errors(res["initial_trans"][i:i+1], g["initial_trans"][i:i+1])
    print(f"pair {i}: dconf {row['max_abs_dconf']:.2e} rot {row['rot_err_deg']:.2e} trans {row['trans_err_m']:.2e} | best seed ours {ours_best} (count {oc}) "
          f"ref {ref_best} (count {rc}); ref count of our best {ref_count_of_ours}; seeds at the ref max count {n_at_max}; "
          f"initial_trans rot {float(ire):.2e} trans {float(ite):.2e}; ref refine {[int(x) for x in g['refine_sizes'][i] if x >= 0]}")
json.dump(rep, open(os.path.join(ROOT, "gpurun_out", f"dump_{case}.json"), "w"), indent=1)
