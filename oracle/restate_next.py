"""CPU restatement of the callers either side of the hot path.  TEST INFRASTRUCTURE
(see oracle/__init__.py): imported by tests/ only.

Pinned by tests/golden/next_*.npz, which oracle/gen_golden_next.py produced by running the UNMODIFIED
reference functions (algorithms/matching.py find_nn / nn_to_mutual, libs/loss.py TransformationLoss /
ClassificationLoss) on seeded inputs.  datasets/LidarFeatureExtractor.py imports open3d and cannot be
imported here; its five matching lines (:109-118) are restated in numpy exactly as written there.
"""
from __future__ import annotations

import numpy as np
import torch


def find_nn(F0, F1, return_2nd=False):
    """algorithms/matching.py:22-65 (the chunking by 250 rows there does not change any value)."""
    dist2 = (F0 ** 2).sum(1).reshape(-1, 1) + (F1 ** 2).sum(1).reshape(1, -1) - 2 * torch.einsum("ac,bc->ab", F0, F1)
    dist = dist2.clamp_min(1e-30).sqrt()
    ind = dist.min(dim=1).indices
    ind2 = None
    if return_2nd:
        d = dist.clone()
        d[torch.arange(len(ind)), ind] = float("inf")
        ind2 = d.min(dim=1).indices
    return torch.arange(len(ind)), ind, ind2, dist


def nn_to_mutual(F0, F1, idx0, idx1):
    """algorithms/matching.py:225-242 with torch_intersect :67-87: pairs present in both directions, ascending i."""
    uniq = torch.unique(idx1)
    _, inv0, _, _ = find_nn(F1[uniq], F0)           # nearest F0 row of every matched F1 row
    back = torch.full((F1.shape[0],), -1, dtype=torch.int64)
    back[uniq] = inv0
    keep = back[idx1] == idx0
    return idx0[keep], idx1[keep]


def lidar_mutual_corr(src_desc, tgt_desc, use_mutual=True):
    """datasets/LidarFeatureExtractor.py:109-118, numpy float32 as there.  Also returns the distance matrix."""
    distance = np.sqrt(2 - 2 * (src_desc @ tgt_desc.T) + 1e-6)
    source_idx = np.argmin(distance, axis=1)
    if use_mutual:
        target_idx = np.argmin(distance, axis=0)
        mutual_nearest = target_idx[source_idx] == np.arange(source_idx.shape[0])
        corr = np.concatenate([np.where(mutual_nearest == 1)[0][:, None], source_idx[mutual_nearest][:, None]], axis=-1)
    else:
        corr = np.concatenate([np.arange(source_idx.shape[0])[:, None], source_idx[:, None]], axis=-1)
    return corr, distance


def transformation_metrics(trans, gt_trans, src_keypts, tgt_keypts):
    """Per pair RE [deg], TE [cm], RMSE, mean squared residual: the loop body of libs/loss.py:44-61 in float64
    (every pair against its own target points)."""
    out = []
    for i in range(trans.shape[0]):
        T, G = trans[i].double(), gt_trans[i].double()
        R, t, Rg, tg = T[:3, :3], T[:3, 3], G[:3, :3], G[:3, 3]
        re = torch.acos(torch.clamp((torch.trace(R.T @ Rg) - 1) / 2.0, min=-1, max=1)) * 180 / np.pi
        te = torch.sqrt(((t - tg) ** 2).sum()) * 100
        warp = src_keypts[i].double() @ R.T + t
        res = warp - tgt_keypts[i].double()
        out.append(torch.stack([re, te, res.norm(dim=-1).mean(), (res ** 2).sum(-1).mean()]))
    return torch.stack(out)


def classification_stats(pred, gt):
    """libs/loss.py:96-112 for one pair without sklearn: counts and the derived precision / recall / f1 /
    mean logits (sklearn returns 0 where a denominator is 0)."""
    pl = pred > 0
    g = gt != 0
    tp = int((pl & g).sum()); fp = int((pl & ~g).sum()); fn = int((~pl & g).sum()); tn = int((~pl & ~g).sum())
    precision = tp / (tp + fp) if tp + fp else 0.0
    recall = tp / (tp + fn) if tp + fn else 0.0
    f1 = 2 * tp / (2 * tp + fp + fn) if 2 * tp + fp + fn else 0.0
    p64, g64 = pred.double(), gt.double()
    return dict(tp=tp, fp=fp, fn=fn, tn=tn, precision=precision, recall=recall, f1=f1,
                logit_true=float((p64 * g64).sum() / max(1.0, float(g64.sum()))),
                logit_false=float((p64 * (1 - g64)).sum() / max(1.0, float((1 - g64).sum()))))
