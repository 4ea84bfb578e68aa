"""Generates tests/golden/next_*.npz by running the UNMODIFIED reference helpers either side of the hot path
(algorithms/matching.py, libs/loss.py) on seeded inputs.  Build container only:  python -m oracle.gen_golden_next
TEST INFRASTRUCTURE.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

from . import synth

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
REF = "/root/reference"


def descriptors(seed, N0, N1, D, noise=0.25):
    """Two descriptor sets sharing min(N0,N1)//2 noisy matches, L2-normalised (like FCGF / FPFH after normalisation)."""
    g = torch.Generator().manual_seed(seed)
    base = torch.randn(max(N0, N1), D, generator=g)
    F0 = base[:N0].clone()
    F1 = torch.randn(N1, D, generator=g)
    m = min(N0, N1) // 2
    perm = torch.randperm(N1, generator=g)[:m]
    F1[perm] = base[:m] + noise * torch.randn(m, D, generator=g)
    F0 = F0 / F0.norm(dim=1, keepdim=True)
    F1 = F1 / F1.norm(dim=1, keepdim=True)
    return F0, F1


def metric_inputs(seed, B, N):
    g = torch.Generator().manual_seed(seed)
    src = torch.rand(B, N, 3, generator=g) * 3
    gt = torch.eye(4).repeat(B, 1, 1)
    est = torch.eye(4).repeat(B, 1, 1)
    for b in range(B):
        R = synth.random_rotation(g)
        gt[b, :3, :3] = R
        gt[b, :3, 3] = torch.randn(3, generator=g)
        # estimate = ground truth perturbed by a rotation of (b+1)*0.7 degrees about a random axis and a shift
        ax = torch.randn(3, generator=g); ax = ax / ax.norm()
        ang = torch.tensor((b + 1) * 0.7 * np.pi / 180)
        K = torch.tensor([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
        dR = torch.eye(3) + torch.sin(ang) * K + (1 - torch.cos(ang)) * (K @ K)
        est[b, :3, :3] = dR @ R
        est[b, :3, 3] = gt[b, :3, 3] + 0.02 * (b + 1) * torch.randn(3, generator=g)
    tgt = torch.einsum("bij,bnj->bni", gt[:, :3, :3], src) + gt[:, None, :3, 3] + 0.01 * torch.randn(B, N, 3, generator=g)
    logits = torch.randn(B, N, generator=g)
    labels = (torch.rand(B, N, generator=g) < 0.3).float()
    logits = logits + 1.5 * (labels - 0.5)
    return src, tgt, est, gt, logits, labels


def main():
    sys.path.insert(0, REF)
    sys.dont_write_bytecode = True
    from algorithms.matching import find_nn, nn_to_mutual                     # the reference, unmodified
    from libs.loss import TransformationLoss, ClassificationLoss
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    torch.set_num_threads(8)
    for name, (seed, N0, N1, D) in {"next_match_fcgf32": (21, 700, 650, 32), "next_match_fpfh33": (22, 500, 900, 33)}.items():
        F0, F1 = descriptors(seed, N0, N1, D)
        i0, i1, i2 = find_nn(F0, F1, return_2nd=True)
        m0, m1 = nn_to_mutual(F0, F1, i0, i1)
        np.savez_compressed(os.path.join(GOLDEN_DIR, name + ".npz"), F0=F0.numpy(), F1=F1.numpy(), idx1=i1.numpy(), idx2=i2.numpy(),
                            mutual0=m0.numpy(), mutual1=m1.numpy())
        print(name, "mutual pairs", len(m0), "of", N0)
    src, tgt, est, gt, logits, labels = metric_inputs(31, 3, 400)
    tl = TransformationLoss(re_thre=15, te_thre=30)
    rows = []
    for b in range(3):      # bs = 1 calls: the reference broadcasts tgt_keypts over the batch (libs/loss.py:53)
        loss, recall, RE, TE, RMSE = tl(est[b:b + 1], gt[b:b + 1], src[b:b + 1], tgt[b:b + 1], logits[b:b + 1])
        rows.append([float(loss), float(recall), float(RE), float(TE), float(RMSE)])
    cl = ClassificationLoss(balanced=True)
    stats = []
    for b in range(3):
        s = cl(logits[b:b + 1], labels[b:b + 1])
        stats.append([s["precision"], s["recall"], s["f1"], s["logit_true"], s["logit_false"]])
    np.savez_compressed(os.path.join(GOLDEN_DIR, "next_metrics.npz"), src=src.numpy(), tgt=tgt.numpy(), est=est.numpy(), gt=gt.numpy(),
                        logits=logits.numpy(), labels=labels.numpy(), trans_loss=np.array(rows, dtype=np.float64),
                        cls_stats=np.array(stats, dtype=np.float64))
    print("next_metrics", rows, stats)


if __name__ == "__main__":
    main()
