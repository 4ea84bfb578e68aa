"""Evaluation metrics on the GPU: the step right after PointDSC.forward.

Host-side mirror of libs/loss.py: TransformationLoss.forward (:18-63) and the statistics half of
ClassificationLoss.forward (:96-112), batched over pairs, one small kernel each, no sklearn / CPU round trip.
There is no CPU path: inputs must be CUDA tensors and the shared library must be built.
"""
from __future__ import annotations

import torch

from . import _cabi


def registration_errors(trans, gt_trans, src_keypts, tgt_keypts):
    """Per pair: fp32 [B,4] = rotation error [deg], translation error [cm], RMSE, mean squared residual."""
    return _cabi.registration_metrics(trans.contiguous().float(), gt_trans.contiguous().float(),
                                      src_keypts.contiguous().float(), tgt_keypts.contiguous().float())


class TransformationLoss(torch.nn.Module):
    """libs/loss.py:11-63.  forward returns (loss, recall [%], RE, TE, RMSE) averaged over the batch, as the reference.
    (The reference broadcasts tgt_keypts over the batch at :53,61, which is only meaningful for bs = 1; here every
    pair is measured against its own target points.)"""

    def __init__(self, re_thre=15, te_thre=30):
        super().__init__()
        self.re_thre = re_thre
        self.te_thre = te_thre

    @torch.no_grad()
    def forward(self, trans, gt_trans, src_keypts, tgt_keypts, probs):
        m = registration_errors(trans, gt_trans, src_keypts, tgt_keypts)
        re, te, rmse, sq = m[:, 0], m[:, 1], m[:, 2], m[:, 3]
        recall = ((te < self.te_thre) & (re < self.re_thre)).float().mean() * 100.0
        has_inlier = (probs > 0).any(dim=1)                       # :57-59: pairs without a predicted inlier add 0
        loss = torch.where(has_inlier, sq, torch.zeros_like(sq)).mean()
        return loss, recall, re.mean(), te.mean(), rmse.mean()


def classification_stats(pred, gt):
    """Per pair precision / recall / f1 (sklearn's zero_division=0 convention) and the mean logit of true / false
    correspondences (libs/loss.py:96-112).  pred: logits or 0/1 labels [B,N]; gt [B,N].  Returns a dict of [B] tensors
    plus the integer counts."""
    counts, sums = _cabi.classification_counts(pred.contiguous().float(), gt.contiguous().float(), 0.0)
    tp, fp, fn, tn = (counts[:, i].double() for i in range(4))
    zero = torch.zeros_like(tp)
    precision = torch.where(tp + fp > 0, tp / (tp + fp).clamp_min(1), zero)
    recall = torch.where(tp + fn > 0, tp / (tp + fn).clamp_min(1), zero)
    f1 = torch.where(2 * tp + fp + fn > 0, 2 * tp / (2 * tp + fp + fn).clamp_min(1), zero)
    n_true, n_false = tp + fn, fp + tn
    return {
        "precision": precision, "recall": recall, "f1": f1,
        "logit_true": sums[:, 0].double() / n_true.clamp_min(1), "logit_false": sums[:, 1].double() / n_false.clamp_min(1),
        "counts": counts,
    }
