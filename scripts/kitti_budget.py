"""Where does the confidence error of the KITTI-shaped 16-pair batch come from?  Per-pair max |dconf| against the fp32
reference and against its float64 evaluation, for several arithmetic variants (tuning library for the chain lengths)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from conftest import load_golden
from oracle import synth
from parity import golden_model
from pointdsc_b200 import _cabi
case = sys.argv[1] if len(sys.argv) > 1 else "batch_kitti_n10000_l12"
meta, g = load_golden(case)
fam = synth.CONFIGS[meta["family"]]
batch = synth.make_batch(meta["pair_seeds"] if "pair_seeds" in meta else [meta["pair_seed"]], meta["N"], **fam)
data = {k: batch[k].cuda() for k in ("corr_pos", "src_keypts", "tgt_keypts")}; data["testing"] = True
ref32 = g["confidence"]
ref64 = g["confidence_f64"] if "confidence_f64" in g else ref32.double()
print("confidence range", float(ref32.min()), float(ref32.max()))
print("reference fp32 vs f64:", " ".join(f"{float(x):.1e}" for x in (ref32.double() - ref64).abs().max(dim=1)[0]))
variants = [("production", {}, {}), ("attn fp32 + mlp fp32", dict(attn_mode=_cabi.ATTN_FP32, mlp_mode="fp32"), {}),
            ("split16 + mlp fp32", dict(attn_mode=_cabi.ATTN_SPLIT16, mlp_mode="fp32"), {}),
            ("production chain=24", {}, {"PDSC_ATTN_CHAIN": "24"}), ("production chain=12", {}, {"PDSC_ATTN_CHAIN": "12"}),
            ("production chain=0 (off)", {}, {"PDSC_ATTN_CHAIN": "0"})]
for name, kw, env in variants:
    if len(sys.argv) > 2 and sys.argv[2] not in name:
        continue
    for k, v in env.items(): os.environ[k] = v
    model, _ = golden_model(meta, **kw)
    if env:
        with _cabi.use_tuning_library():
            conf = model(data)["confidence"].cpu()
    else:
        conf = model(data)["confidence"].cpu()
    for k in env: del os.environ[k]
    d32 = (conf - ref32).abs().max(dim=1)[0]; d64 = (conf.double() - ref64).abs().max(dim=1)[0]
    print(f"{name:28s} vs fp32 ref: max {float(d32.max()):.2e} | " + " ".join(f"{float(x):.1e}" for x in d32))
    print(f"{'':28s} vs f64  ref: max {float(d64.max()):.2e} | " + " ".join(f"{float(x):.1e}" for x in d64), flush=True)
