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
}


# BASELINE.json's full sizes (configs[1] and configs[2]), 12 layers: lean fixtures (outputs only; the inputs regenerate
# from the seeds), written by `python -m oracle.gen_golden --full` (the N=10000 run needs ~6 GB and a minute of CPU)
FULL_CASES = {
    # weight seed 99: all confidences positive, so every seed key is positive and the seed order is well defined (with seed
    # 1234 every confidence is negative at 12 layers and the reference's argsort fills the seed list from exact ties)
    "full_3dmatch_n5000_l12": ("3dmatch", 5000, 12, 0, 99),
    "full_kitti_n10000_l12": ("kitti", 10000, 12, 100, 99),
}


def case_inputs(name):
    fam, N, L, pair_seed, w_seed = (CASES.get(name) or FULL_CASES[name])
    cfg = synth.CONFIGS[fam]
    sd = synth.make_state_dict(w_seed, num_layers=L, sigma_d=cfg["sigma_d"])
    pair = synth.make_pair(pair_seed, N, **cfg)
    ctor = dict(num_layers=L, inlier_threshold=cfg["inlier_threshold"], sigma_d=cfg["sigma_d"],
                nms_radius=cfg["nms_radius"])
    return cfg, sd, pair, ctor


def main():
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    torch.set_num_threads(8)
    for name in CASES:
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
        }
        meta = dict(case=name, family=CASES[name][0], N=CASES[name][1], num_layers=CASES[name][2],
                    pair_seed=CASES[name][3], weight_seed=CASES[name][4], ctor=ctor,
                    torch=torch.__version__, reference="AmnonDrory/PointDSC models/PointDSC.py forward (testing)")
        np.savez_compressed(os.path.join(GOLDEN_DIR, name + ".npz"),
                            meta=np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8),
                            **{k: v.numpy() for k, v in arrays.items()})
        print(name, "labels", int(rec["final_labels"].sum()), "refine", rec["refine_sizes"],
              "conf [%.4f, %.4f]" % (rec["confidence"].min(), rec["confidence"].max()))


def main_full():
    torch.set_num_threads(os.cpu_count())
    for name in FULL_CASES:
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
        }
        meta = dict(case=name, family=FULL_CASES[name][0], N=FULL_CASES[name][1], num_layers=FULL_CASES[name][2],
                    pair_seed=FULL_CASES[name][3], weight_seed=FULL_CASES[name][4], ctor=ctor, k=k,
                    torch=torch.__version__, reference="AmnonDrory/PointDSC models/PointDSC.py forward (testing)")
        np.savez_compressed(os.path.join(GOLDEN_DIR, name + ".npz"),
                            meta=np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8),
                            **{k_: v.numpy() for k_, v in arrays.items()})
        print(name, "labels", int(rec["final_labels"].sum()), "refine", rec["refine_sizes"],
              "conf [%.4f, %.4f]" % (rec["confidence"].min(), rec["confidence"].max()),
              os.path.getsize(os.path.join(GOLDEN_DIR, name + ".npz")) // 1024, "KiB")


if __name__ == "__main__":
    import sys
    main_full() if "--full" in sys.argv else main()
