"""Putative-correspondence construction on the GPU: the step right before PointDSC.forward.

Host-side mirror of the reference helpers, same names and argument meaning, CUDA tensors in and out
(the reference returns CPU tensors / numpy arrays; nothing here touches the host):

  find_nn, nn_to_mutual              algorithms/matching.py:22-65, 225-242 (+ torch_intersect :67-87)
  mutual_nn_correspondences          datasets/LidarFeatureExtractor.py:109-118
  correspondence_labels              datasets/LidarFeatureExtractor.py:120-124

There is no CPU path: inputs must be CUDA tensors and the shared library must be built.
"""
from __future__ import annotations

import torch

from . import _cabi


def _f32(x):
    return x.contiguous().float()


def find_nn(F0, F1, return_2nd=False):
    """For every row of F0 [N0,D] the nearest row of F1 [N1,D] in L2 (algorithms/matching.py:22-65).
    Returns (corres_idx0 = arange(N0), corres_idx1, idx1_2nd or None), int64 on F0's device."""
    idx1, idx2, _, _ = _cabi.nn_match(_f32(F0)[None], _f32(F1)[None], metric=_cabi.MATCH_L2, want_second=return_2nd)
    idx0 = torch.arange(F0.shape[0], dtype=torch.int64, device=F0.device)
    return idx0, idx1[0], (idx2[0] if return_2nd else None)


def nn_to_mutual(feats0, feats1, corres_idx0, corres_idx1, idx1_2nd=None, force_return_2nd=False):
    """Keeps the pairs (i, j) with j = nn(i) and i = nn(j) (algorithms/matching.py:225-242), in ascending i.
    corres_idx0 must be the full range [0, N0), as the reference itself relies on (:238)."""
    N0, N1 = feats0.shape[0], feats1.shape[0]
    if corres_idx0.numel() != N0:
        raise ValueError("nn_to_mutual: corres_idx0 must be arange(len(feats0)) (algorithms/matching.py:238)")
    idx_ba, _, _, _ = _cabi.nn_match(_f32(feats1)[None], _f32(feats0)[None], metric=_cabi.MATCH_L2)
    corr, count = _cabi.mutual_corr(corres_idx1.to(feats0.device).contiguous()[None], idx_ba, N1, use_mutual=True)
    m = int(count[0])                                   # the one host read: the size of the result
    final0, final1 = corr[0, :m, 0], corr[0, :m, 1]
    if idx1_2nd is not None:
        return final0, final1, idx1_2nd.to(feats0.device)[final0]
    if force_return_2nd:
        return final0, final1, None
    return final0, final1


def mutual_nn_correspondences(src_desc, tgt_desc, use_mutual=True):
    """corr [M,2] int64 from L2-normalised descriptors (datasets/LidarFeatureExtractor.py:109-118): source i is paired
    with argmin_j sqrt(2 - 2 s_i.t_j + 1e-6); with use_mutual only pairs that are also the argmin over sources stay."""
    s, t = _f32(src_desc)[None], _f32(tgt_desc)[None]
    idx_ab, _, _, _ = _cabi.nn_match(s, t, metric=_cabi.MATCH_DOT)
    idx_ba = _cabi.nn_match(t, s, metric=_cabi.MATCH_DOT)[0] if use_mutual else None
    corr, count = _cabi.mutual_corr(idx_ab, idx_ba, t.shape[1], use_mutual=use_mutual)
    return corr[0, : int(count[0])]


def correspondence_labels(src_keypts, tgt_keypts, corr, gt_trans, inlier_threshold):
    """labels [M] (int64 0/1): |T_gt src[corr0] - tgt[corr1]| < inlier_threshold (LidarFeatureExtractor.py:120-124)."""
    a = src_keypts[corr[:, 0]].float()
    b = tgt_keypts[corr[:, 1]].float()
    T = gt_trans.to(a.device).float()
    warped = a @ T[:3, :3].T + T[:3, 3]
    return ((warped - b).pow(2).sum(-1).sqrt() < inlier_threshold).long()
