"""The optional ICP hand-off after PointDSC.forward, on the GPU.

Host-side mirror of `evaluation/benchmark_utils.py:40-56 icp_refine`, same name and argument meaning; the reference
moves the key points to the host and calls open3d's point-to-point `registration_icp` with a 0.1 m correspondence
distance and the default convergence criteria, one pair at a time.  Here B pairs are refined in one call on the device
(no CPU path).  `registration_ransac_based_on_correspondence` (evaluation/test_3DMatch.py:59-80) draws from open3d's own
random stream and is not reproduced.
"""
from __future__ import annotations

import torch

from . import _cabi


def icp_refine(src_keypts, tgt_keypts, pred_trans, max_correspondence_distance=0.1, max_iteration=30,
               relative_fitness=1e-6, relative_rmse=1e-6, return_stats=False):
    """src_keypts, tgt_keypts [B,N,3], pred_trans [B,4,4] (CUDA) -> refined_T [B,4,4] fp32 on the same device."""
    out, stats, iters, _ = _cabi.icp_refine(src_keypts.float().contiguous(), tgt_keypts.float().contiguous(),
                                            pred_trans.float().contiguous(), max_correspondence_distance, max_iteration,
                                            relative_fitness, relative_rmse)
    if return_stats:
        return out, {"fitness": stats[:, 0], "inlier_rmse": stats[:, 1], "iterations": iters}
    return out
