"""Generates ``tests/golden/*.npz`` by running the UNMODIFIED reference
(``/root/reference/models/PointDSC.py``) on seeded synthetic inputs.

Run in the build container only:  ``python -m oracle.gen_golden``
The fixtures pin ``oracle/restate.py`` (tests/test_oracle.py) and are the
reference-produced targets of the GPU parity tests.  TEST INFRASTRUCTURE.
"""
from __future__ import annotations

import json
import os

import numpy as np
import torch

from . import ref_harness, synth

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

# name -> (config family, N, num_layers, pair seed, weight seed, conf_bias)
CASES = {
    "g0_3dmatch_n256_l2": ("3dmatch", 256, 2, 11, 1234),
    "g1_3dmatch_n1000_l12": ("3dmatch", 1000, 12, 7, 1234),
    "g2_kitti_n800_l6": ("kitti", 800, 6, 107, 4321),
    "g3_3dmatch_n2000_l12": ("3dmatch", 2000, 12, 3, 99),
    # stress weights (synth.STRESS): sharp attention (|logit| up to ~20 in log2 units), confidences spanning > 1
    "g4_3dmatch_n2000_l12_stress": ("3dmatch", 2000, 12, 5, 99, synth.STRESS),
    # 40 duplicated / near-duplicated inlier correspondences (synth.make_pair(duplicates=40)): the seed kNN's rank 0 is not
    # always the seed itself (7th field = extra make_pair arguments)
    "g5_3dmatch_n1000_l12_dups": ("3dmatch", 1000, 12, 21, 99, {}, dict(duplicates=40)),
}


# BASELINE.json's full sizes (configs[1] and configs[2]), 12 layers: lean fixtures (outputs only; the inputs regenerate
# from the seeds), written by `python -m oracle.gen_golden --full` (the N=10000 run needs ~6 GB and a minute of CPU)
FULL_CASES = {
    # weight seed 99: all confidences positive, so every seed key is positive and the seed order is well defined (with seed
    # 1234 every confidence is negative at 12 layers and the reference's argsort fills the seed list from exact ties)
    "full_3dmatch_n5000_l12": ("3dmatch", 5000, 12, 0, 99),
    "full_kitti_n10000_l12": ("kitti", 10000, 12, 100, 99),
    "full_3dmatch_n5000_l12_stress": ("3dmatch", 5000, 12, 0, 99, synth.STRESS),
}


def weight_kwargs(name):
    c = CASES.get(name) or FULL_CASES[name]
    return dict(c[5]) if len(c) > 5 else {}


def pair_kwargs(name):
    c = CASES.get(name) or FULL_CASES[name]
    return dict(c[6]) if len(c) > 6 else {}


def case_inputs(name):
    fam, N, L, pair_seed, w_seed = (CASES.get(name) or FULL_CASES[name])[:5]
    cfg = synth.CONFIGS[fam]
    sd = synth.make_state_dict(w_seed, num_layers=L, sigma_d=cfg["sigma_d"], **weight_kwargs(name))
    pair = synth.make_pair(pair_seed, N, **cfg, **pair_kwargs(name))
    ctor = dict(num_layers=L, inlier_threshold=cfg["inlier_threshold"], sigma_d=cfg["sigma_d"],
                nms_radius=cfg["nms_radius"])
    return cfg, sd, pair, ctor


def main(only=None):
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    torch.set_num_threads(8)
    for name in CASES:
        if only and only not in name:
            continue
        cfg, sd, pair, ctor = case_inputs(name)
        model = ref_harness.build_reference_model(sd, **ctor)
        rec = ref_harness.run_reference(model, pair["corr_pos"], pair["src_keypts"], pair["tgt_keypts"])
        arrays = {
            "corr_pos": pair["corr_pos"], "src_keypts": pair["src_keypts"], "tgt_keypts": pair["tgt_keypts"],
            "confidence": rec["confidence"], "seeds": rec["seeds"], "knn_idx": rec["knn_idx"],
            "eigvec": rec["eigvec"], "seed_trans": rec["seed_trans"], "seed_fitness": rec["seed_fitness"],
            "initial_trans": rec["initial_trans"], "final_labels": rec["final_labels"],
            "final_trans": rec["final_trans"],
            "refine_sizes": torch.tensor(rec["refine_sizes"], dtype=torch.int64),
            "normed_features": rec["normed_features"],
            "power_iters": torch.tensor([rec["power_iters"]], dtype=torch.int64),
        }
        meta = dict(case=name, family=CASES[name][0], N=CASES[name][1], num_layers=CASES[name][2],
                    pair_seed=CASES[name][3], weight_seed=CASES[name][4], ctor=ctor, weight_kwargs=weight_kwargs(name),
                    pair_kwargs=pair_kwargs(name), torch=torch.__version__, reference="AmnonDrory/PointDSC models/PointDSC.py forward (testing)")
        np.savez_compressed(os.path.join(GOLDEN_DIR, name + ".npz"),
                            meta=np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8),
                            **{k: v.numpy() for k, v in arrays.items()})
        print(name, "labels", int(rec["final_labels"].sum()), "refine", rec["refine_sizes"],
              "conf [%.4f, %.4f]" % (rec["confidence"].min(), rec["confidence"].max()))


def main_full(only=None):
    torch.set_num_threads(os.cpu_count())
    for name in FULL_CASES:
        if only and only not in name:
            continue
        cfg, sd, pair, ctor = case_inputs(name)
        model = ref_harness.build_reference_model(sd, **ctor)
        rec = ref_harness.run_reference(model, pair["corr_pos"], pair["src_keypts"], pair["tgt_keypts"])
        k = rec["knn_idx"].shape[-1]
        # the same unmodified reference in float64: how far its own fp32 result is from exact arithmetic (the noise floor any
        # fp32 implementation with another summation order lives on; 3.7e-5 in the confidences at KITTI scale)
        torch.set_default_dtype(torch.float64)
        try:
            rec64 = ref_harness.run_reference(ref_harness.build_reference_model(sd, **ctor).double(), pair["corr_pos"].double(),
                                              pair["src_keypts"].double(), pair["tgt_keypts"].double())
        finally:
            torch.set_default_dtype(torch.float32)
        arrays = {
            "confidence_f64": rec64["confidence"], "final_trans_f64": rec64["final_trans"],
            "confidence": rec["confidence"], "seeds": rec["seeds"].to(torch.int32), "knn_idx": rec["knn_idx"].to(torch.int32),
            "seed_trans": rec["seed_trans"], "inlier_count": torch.round(rec["seed_fitness"] * FULL_CASES[name][1]).to(torch.int32),
            "initial_trans": rec["initial_trans"], "final_labels": rec["final_labels"].to(torch.uint8),
            "final_trans": rec["final_trans"], "refine_sizes": torch.tensor(rec["refine_sizes"], dtype=torch.int64),
            "power_iters": torch.tensor([rec["power_iters"]], dtype=torch.int64),
        }
        meta = dict(case=name, family=FULL_CASES[name][0], N=FULL_CASES[name][1], num_layers=FULL_CASES[name][2],
                    pair_seed=FULL_CASES[name][3], weight_seed=FULL_CASES[name][4], ctor=ctor, k=k,
                    weight_kwargs=weight_kwargs(name), torch=torch.__version__, reference="AmnonDrory/PointDSC models/PointDSC.py forward (testing)")
        np.savez_compressed(os.path.join(GOLDEN_DIR, name + ".npz"),
                            meta=np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8),
                            **{k_: v.numpy() for k_, v in arrays.items()})
        print(name, "labels", int(rec["final_labels"].sum()), "refine", rec["refine_sizes"],
              "conf [%.4f, %.4f]" % (rec["confidence"].min(), rec["confidence"].max()),
              os.path.getsize(os.path.join(GOLDEN_DIR, name + ".npz")) // 1024, "KiB")


# The whole timed batches of bench.py / BASELINE configs[1] and configs[2]: 16 pairs each (pair seeds 0..15 and 100..115,
# weight seed 99), every pair through the unmodified reference one at a time (its test path is batch-size-1 only).  Lean:
# outputs only, stacked over the pairs; written by `python -m oracle.gen_golden --batch`.
BATCH_CASES = {
    "batch_3dmatch_n5000_l12": ("3dmatch", 5000, 12, list(range(0, 16)), 99),
    "batch_kitti_n10000_l12": ("kitti", 10000, 12, list(range(100, 116)), 99),
}


def main_batch(only=None, reuse_f64=False):
    """reuse_f64: keep the float64 columns of the existing fixture (they come from this same function; the float64 pass is
    the slow part) and redo the fp32 pass only."""
    torch.set_num_threads(os.cpu_count())
    for name, (fam, N, L, pair_seeds, w_seed) in BATCH_CASES.items():
        if only and only not in name:
            continue
        cfg = synth.CONFIGS[fam]
        sd = synth.make_state_dict(w_seed, num_layers=L, sigma_d=cfg["sigma_d"])
        ctor = dict(num_layers=L, inlier_threshold=cfg["inlier_threshold"], sigma_d=cfg["sigma_d"], nms_radius=cfg["nms_radius"])
        model = ref_harness.build_reference_model(sd, **ctor)
        cols = {k: [] for k in ("confidence", "seeds", "inlier_count", "initial_trans", "final_labels", "final_trans",
                                "refine_sizes", "power_iters", "best_idx", "best_knn_idx", "confidence_f64", "final_trans_f64")}
        old = None
        if reuse_f64 and os.path.exists(os.path.join(GOLDEN_DIR, name + ".npz")):
            old = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
            old = old if "confidence_f64" in old.files else None
        torch.set_default_dtype(torch.float64)
        try:
            model64 = ref_harness.build_reference_model(sd, **ctor).double()
        finally:
            torch.set_default_dtype(torch.float32)
        for pi, ps in enumerate(pair_seeds):
            pair = synth.make_pair(ps, N, **cfg)
            rec = ref_harness.run_reference(model, pair["corr_pos"], pair["src_keypts"], pair["tgt_keypts"])
            counts = torch.round(rec["seed_fitness"] * N).to(torch.int32)
            # the neighbour list of the winning seed (models/PointDSC.py:251-252): where a near-tie in feature distance gives
            # another list, the tests re-run that one hypothesis on the reference's list (tests/parity.py)
            cols["best_knn_idx"].append(rec["knn_idx"][0, int(torch.argmax(rec["seed_fitness"][0]))].to(torch.int32))
            sizes = rec["refine_sizes"] + [-1] * (20 - len(rec["refine_sizes"]))
            cols["confidence"].append(rec["confidence"][0])
            cols["seeds"].append(rec["seeds"][0].to(torch.int32))
            cols["inlier_count"].append(counts[0])
            cols["best_idx"].append(torch.argmax(rec["seed_fitness"][0]).to(torch.int32))
            cols["initial_trans"].append(rec["initial_trans"][0])
            cols["final_labels"].append(rec["final_labels"][0].to(torch.uint8))
            cols["final_trans"].append(rec["final_trans"][0])
            cols["refine_sizes"].append(torch.tensor(sizes, dtype=torch.int32))
            cols["power_iters"].append(torch.tensor(rec["power_iters"], dtype=torch.int32))
            # the same unmodified reference in float64: this pair's fp32 noise floor (how far the reference's own fp32 result
            # is from exact arithmetic)
            if old is not None:
                cols["confidence_f64"].append(torch.from_numpy(old["confidence_f64"][pi]))
                cols["final_trans_f64"].append(torch.from_numpy(old["final_trans_f64"][pi]))
                continue
            torch.set_default_dtype(torch.float64)
            try:
                rec64 = ref_harness.run_reference(model64, pair["corr_pos"].double(), pair["src_keypts"].double(),
                                                  pair["tgt_keypts"].double())
            finally:
                torch.set_default_dtype(torch.float32)
            cols["confidence_f64"].append(rec64["confidence"][0])
            cols["final_trans_f64"].append(rec64["final_trans"][0])
            print("   fp32 reference vs float64 reference: max |dconf|", float((rec["confidence"][0].double() - rec64["confidence"][0]).abs().max()), flush=True)
            print(name, "pair seed", ps, "labels", int(rec["final_labels"].sum()), "power iters", rec["power_iters"],
                  "refine", rec["refine_sizes"], flush=True)
        arrays = {k: torch.stack(v) for k, v in cols.items()}
        meta = dict(case=name, family=fam, N=N, num_layers=L, pair_seeds=pair_seeds, weight_seed=w_seed, ctor=ctor,
                    torch=torch.__version__, reference="AmnonDrory/PointDSC models/PointDSC.py forward (testing), one pair at a time")
        path = os.path.join(GOLDEN_DIR, name + ".npz")
        np.savez_compressed(path, meta=np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8),
                            **{k_: v.numpy() for k_, v in arrays.items()})
        print(name, os.path.getsize(path) // 1024, "KiB", flush=True)


if __name__ == "__main__":
    import sys
    flags = [a for a in sys.argv[1:] if a.startswith("--")]
    only = next((a for a in sys.argv[1:] if not a.startswith("--")), None)       # substring filter on the case names
    if "--batch" in flags:
        main_batch(only, reuse_f64="--reuse-f64" in flags)
    elif "--full" in flags:
        main_full(only)
    else:
        main(only)
