"""Confidence error of the forward at BASELINE's full sizes, per precision mode and batch size, against the unmodified
reference in fp32 and in float64 (tests/golden/full_*.npz).  python scripts/error_budget.py > profiles/r01_error_budget.json"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from conftest import load_golden
from oracle import synth
from parity import golden_model, trans_errors
from pointdsc_b200 import _cabi

out = {}
for case in ("full_3dmatch_n5000_l12", "full_kitti_n10000_l12"):
    meta, g = load_golden(case)
    cfg = synth.CONFIGS[meta["family"]]
    rows = []
    for attn, mlp in (("fp32", "fp32"), ("split16", "fp32"), ("split16", "split16"), ("split16", "fused")):
        for B in (1, 16):
            model, _ = golden_model(meta, attn_mode={"fp32": _cabi.ATTN_FP32, "split16": _cabi.ATTN_SPLIT16}[attn], mlp_mode=mlp)
            batch = synth.make_batch([meta["pair_seed"] + i for i in range(B)], meta["N"], **cfg)
            data = {k: batch[k].cuda() for k in ("corr_pos", "src_keypts", "tgt_keypts")}
            data["testing"] = True
            res = model(data)
            c = res["confidence"][0:1].cpu()
            re, te = trans_errors(res["final_trans"][0:1], g["final_trans"])
            rows.append({"attention": attn, "mlp": mlp, "batch": B,
                         "dconf_vs_fp32_reference": float((c - g["confidence"]).abs().max()),
                         "dconf_vs_float64_reference": float((c.double() - g["confidence_f64"]).abs().max()),
                         "rms_vs_float64_reference": float((c.double() - g["confidence_f64"]).pow(2).mean().sqrt()),
                         "rot_err_deg": float(re.max()), "trans_err_m": float(te.max()),
                         "label_mismatches": int((res["final_labels"][0].cpu() != g["final_labels"][0].float()).sum())})
    out[case] = {"fp32_reference_vs_float64_reference": float((g["confidence"].double() - g["confidence_f64"]).abs().max()),
                 "fp32_reference_rms_vs_float64": float((g["confidence"].double() - g["confidence_f64"]).pow(2).mean().sqrt()),
                 "confidence_range": [float(g["confidence"].min()), float(g["confidence"].max())], "runs": rows}
print(json.dumps(out, indent=1))
