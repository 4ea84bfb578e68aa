"""CPU restatement of the ICP hand-off: `evaluation/benchmark_utils.py:40-56 icp_refine`.  TEST
INFRASTRUCTURE (see oracle/__init__.py).

**Parity unpinned**: the reference calls open3d's `registration_icp` (pinned open3d 0.9.0.0, `environment.yml:76`;
not vendored, not installed here).  This restates that routine's published algorithm (Besl & McKay point-to-point ICP
as open3d implements it) in float64:

    result = associate(T0)                               # nearest target within max_dist for every transformed source point
    repeat max_iteration times:
        update = umeyama(pairs, no scaling)              # identity when there are no pairs
        T = update @ T
        previous, result = result, associate(T)
        stop if |d fitness| < relative_fitness and |d rmse| < relative_rmse

fitness = pairs / N, inlier_rmse = sqrt(sum d^2 / pairs).  Nearest = smallest (dist^2, target index) with
dist^2 < max_dist^2, dist^2 = dx*dx + dy*dy + dz*dz (the tie rule is ours; FLANN leaves it open).
"""
from __future__ import annotations

import numpy as np


def _transform(T, P):
    return np.stack([T[r, 0] * P[:, 0] + T[r, 1] * P[:, 1] + T[r, 2] * P[:, 2] + T[r, 3] for r in range(3)], 1)


def associate(P, tgt, tree, max_dist):
    """idx [N] (-1 = none) and dist^2 [N] of the nearest target point closer than max_dist."""
    n = len(tgt)
    k = min(4, n)
    _, j = tree.query(P, k=k)
    j = j.reshape(len(P), k)
    valid = j < n
    q = tgt[np.where(valid, j, 0)]
    d = q - P[:, None, :]
    d2 = d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1] + d[..., 2] * d[..., 2]
    d2 = np.where(valid & (d2 < max_dist * max_dist), d2, np.inf)
    order = np.lexsort((j, d2), axis=1)[:, 0]
    rows = np.arange(len(P))
    best_d2 = d2[rows, order]
    idx = np.where(np.isfinite(best_d2), j[rows, order], -1)
    return idx, best_d2


def umeyama_rigid(P, Q):
    """R, t minimising sum |R p + t - q|^2 (no scaling); identity for no points."""
    if len(P) == 0:
        return np.eye(3), np.zeros(3)
    mp, mq = P.mean(0), Q.mean(0)
    H = (P - mp).T @ (Q - mq) / len(P)
    U, _, Vt = np.linalg.svd(H)
    V = Vt.T
    D = np.diag([1.0, 1.0, np.linalg.det(V @ U.T)])
    R = V @ D @ U.T
    return R, mq - R @ mp


def icp_point_to_point(src, tgt, T0, max_dist=0.1, max_iter=30, rel_fitness=1e-6, rel_rmse=1e-6):
    """One pair.  Returns (T [4,4] float64, fitness, inlier_rmse, updates applied, final idx)."""
    from scipy.spatial import cKDTree
    src, tgt = np.asarray(src, np.float64), np.asarray(tgt, np.float64)
    T = np.asarray(T0, np.float64).copy()
    tree = cKDTree(tgt)

    def stats(idx, d2):
        m = idx >= 0
        n = int(m.sum())
        return n / len(src), (np.sqrt(d2[m].sum() / n) if n else 0.0)

    idx, d2 = associate(_transform(T, src), tgt, tree, max_dist)
    fit, rmse = stats(idx, d2)
    updates = 0
    for _ in range(max_iter):
        m = idx >= 0
        R, t = umeyama_rigid(_transform(T, src)[m], tgt[idx[m]])
        U = np.eye(4); U[:3, :3] = R; U[:3, 3] = t
        T = U @ T
        updates += 1
        idx, d2 = associate(_transform(T, src), tgt, tree, max_dist)
        nfit, nrmse = stats(idx, d2)
        stop = abs(fit - nfit) < rel_fitness and abs(rmse - nrmse) < rel_rmse
        fit, rmse = nfit, nrmse
        if stop:
            break
    return T, fit, rmse, updates, idx
