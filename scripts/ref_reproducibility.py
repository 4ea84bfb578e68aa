"""How reproducible is the reference's OWN fp32 result?  The same pair through oracle/restate.py (bit-identical to the
unmodified reference, tests/test_oracle.py) with the correspondences in another order and with another thread count: the
mathematics is unchanged, only the fp32 summation order moves.  CPU only.
python scripts/ref_reproducibility.py kitti 10000 105 [stress]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from oracle import restate, synth
fam, N, seed = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
kw = synth.STRESS if len(sys.argv) > 4 and sys.argv[4] == "stress" else {}
cfg = synth.CONFIGS[fam]
sd = synth.make_state_dict(99, num_layers=12, sigma_d=cfg["sigma_d"], **kw)
p = synth.make_pair(seed, N, **cfg)

def conf(pair, threads):
    torch.set_num_threads(threads)
    _, compat = restate.spatial_compatibility(pair["src_keypts"], pair["tgt_keypts"], sd["sigma_spat"])
    return restate.classification(sd, restate.encoder(sd, pair["corr_pos"], compat, 12))[0]

base = conf(p, os.cpu_count())
out = {"case": f"{fam} N={N} pair seed {seed}" + (" stress weights" if kw else ""), "conf_range": [float(base.min()), float(base.max())]}
g = torch.Generator().manual_seed(1)
for name in ("permuted_a", "permuted_b"):
    perm = torch.randperm(N, generator=g)
    q = {k: p[k][:, perm].contiguous() for k in ("corr_pos", "src_keypts", "tgt_keypts")}
    c = conf(q, os.cpu_count())
    inv = torch.empty_like(perm); inv[perm] = torch.arange(N)
    out[name] = float((c[inv] - base).abs().max())
out["one_thread"] = float((conf(p, 1) - base).abs().max())
print(json.dumps(out), flush=True)
