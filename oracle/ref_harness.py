"""Runs the REAL reference (``/root/reference/models/PointDSC.py``, imported
read-only) and records its intermediates.  TEST INFRASTRUCTURE; only usable in
the build container: ``/root/reference`` does not exist on the GPU box, so nothing
in ``-m gpu`` tests, ``smoke()`` or ``bench.py`` imports this module.
"""
from __future__ import annotations

import os
import sys

import torch

REF_ROOT = os.environ.get("POINTDSC_REFERENCE", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "models", "PointDSC.py"))


def _import_reference():
    sys.dont_write_bytecode = True            # the tree is read-only
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    import models.PointDSC as ref_mod         # noqa: E402
    import models.common as ref_common        # noqa: E402
    return ref_mod, ref_common


def build_reference_model(state_dict, **ctor):
    ref_mod, _ = _import_reference()
    m = ref_mod.PointDSC(**ctor)
    m.load_state_dict(state_dict, strict=True)
    return m.eval()


@torch.no_grad()
def run_reference(model, corr_pos, src, tgt):
    """One pair through the unmodified reference ``forward``; intermediates are captured
    by wrapping bound methods / module-level helpers."""
    ref_mod, ref_common = _import_reference()
    rec = {}

    orig_pick = model.pick_seeds
    orig_cst = model.cal_seed_trans
    orig_eig = model.cal_leading_eigenvector
    orig_cls = model.classification.forward
    orig_knn = ref_mod.knn
    orig_rt = ref_mod.rigid_transform_3d

    def pick(dists, scores, R, max_num):
        s = orig_pick(dists, scores, R, max_num)
        rec["seeds"] = s.clone()
        rec["confidence"] = scores.clone()
        return s

    def cst(seeds, feats, s, t):
        rec["normed_features"] = feats.clone()
        r = orig_cst(seeds, feats, s, t)
        rec["seed_trans"], rec["seed_fitness"], rec["initial_trans"], rec["final_labels"] = \
            [x.clone() for x in r]
        return r

    def eig(M, method="power"):
        rec["local_M"] = M.clone()
        # the loop (models/PointDSC.py:351-356) calls torch.allclose exactly once per iteration it runs: counting those calls
        # records the integer side effect "iterations used" without touching the reference
        orig_allclose = torch.allclose
        calls = []

        def counting_allclose(*a, **kw):
            calls.append(1)
            return orig_allclose(*a, **kw)

        torch.allclose = counting_allclose
        try:
            v = orig_eig(M, method)
        finally:
            torch.allclose = orig_allclose
        rec["eigvec"] = v.clone()
        rec["power_iters"] = len(calls)
        return v

    def knn(x, k, ignore_self=False, normalized=True):
        idx = orig_knn(x, k, ignore_self=ignore_self, normalized=normalized)
        rec["knn_idx_all"] = idx.clone()
        return idx

    rt_calls = []

    def rt(A, B, weights=None, weight_threshold=0):
        T = orig_rt(A, B, weights, weight_threshold)
        rt_calls.append(dict(n=A.shape[1], T=T.clone()))
        return T

    def cls(x):
        rec["corr_features"] = x.permute(0, 2, 1).clone()
        return orig_cls(x)

    model.pick_seeds, model.cal_seed_trans, model.cal_leading_eigenvector = pick, cst, eig
    model.classification.forward = cls
    ref_mod.knn, ref_mod.rigid_transform_3d = knn, rt
    try:
        res = model({"corr_pos": corr_pos, "src_keypts": src, "tgt_keypts": tgt, "testing": True})
    finally:
        model.pick_seeds, model.cal_seed_trans, model.cal_leading_eigenvector = orig_pick, orig_cst, orig_eig
        model.classification.forward = orig_cls
        ref_mod.knn, ref_mod.rigid_transform_3d = orig_knn, orig_rt
    rec["final_trans"] = res["final_trans"].clone()
    assert res["M"] is None
    assert torch.equal(res["final_labels"], rec["final_labels"])
    k = rec["local_M"].shape[-1]
    rec["knn_idx"] = rec["knn_idx_all"].gather(1, rec["seeds"][:, :, None].expand(-1, -1, k))
    # first rigid_transform_3d call is the per-seed one; the rest are post-refinement steps
    rec["refine_sizes"] = [c["n"] for c in rt_calls[1:]]
    rec["refine_trans"] = [c["T"] for c in rt_calls[1:]]
    return rec
