"""CPU restatement of the reference hot path, stage by stage, with every
intermediate exposed.  TEST INFRASTRUCTURE (see oracle/__init__.py).

Each function cites the reference lines it follows (paths relative to the
reference checkout).  It works on ONE pair at a time (the reference's test
path is batch-size-1 only: ``models/PointDSC.py:210,414``) and on a plain
``state_dict`` instead of an ``nn.Module``.  The tensor ops are kept in the same
order and dtype as the reference so that, on the same torch build, the results
are bit-identical to the reference's (checked by ``tests/test_oracle.py`` against
the golden vectors produced by the real reference, ``oracle/gen_golden.py``).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

BN_EPS = 1e-5  # nn.BatchNorm1d default, models/PointDSC.py:14,59


# --------------------------------------------------------------------------
# small helpers
# --------------------------------------------------------------------------
def _conv(sd, name, x):
    """1x1 Conv1d on [1,C,N] (models/PointDSC.py:13,21-23,54,57,108-112)."""
    return F.conv1d(x, sd[name + ".weight"], sd[name + ".bias"])


def _bn(sd, name, x):
    """BatchNorm1d in eval mode (running stats)."""
    return F.batch_norm(x, sd[name + ".running_mean"], sd[name + ".running_var"],
                        sd[name + ".weight"], sd[name + ".bias"], False, 0.0, BN_EPS)


def integrate_trans(R, t):
    """utils/SE3.py:73-96 (batched branch)."""
    T = torch.eye(4)[None].repeat(R.shape[0], 1, 1)
    T[:, :3, :3] = R
    T[:, :3, 3:4] = t.view(-1, 3, 1)
    return T


def transform(pts, T):
    """utils/SE3.py:43-57 (batched branch): R @ pts^T + t."""
    return (T[:, :3, :3] @ pts.permute(0, 2, 1) + T[:, :3, 3:4]).permute(0, 2, 1)


# --------------------------------------------------------------------------
# R1  pairwise distances and spatial compatibility  (models/PointDSC.py:150-153)
# --------------------------------------------------------------------------
def spatial_compatibility(src, tgt, sigma_spat):
    """src,tgt [1,N,3] -> src_dist [1,N,N], compat [1,N,N]."""
    src_dist = torch.norm(src[:, :, None, :] - src[:, None, :, :], dim=-1)
    c = src_dist - torch.norm(tgt[:, :, None, :] - tgt[:, None, :, :], dim=-1)
    compat = torch.clamp(1.0 - c ** 2 / sigma_spat ** 2, min=0)
    return src_dist, compat


# --------------------------------------------------------------------------
# R3  compatibility-weighted non-local attention  (models/PointDSC.py:27-45)
# --------------------------------------------------------------------------
def nonlocal_message(q, k, v, compat):
    """q,k,v [1,C,N], compat [1,N,N] -> message [1,C,N] (lines 36-42, one head)."""
    C = q.shape[1]
    logits = torch.einsum("bco,bci->boi", q, k) / C ** 0.5
    w = torch.softmax(compat * logits, dim=-1)
    return torch.einsum("boi,bci->bco", w, v)


def nonlocal_block(sd, prefix, feat, compat, taps=None):
    q = _conv(sd, prefix + ".projection_q", feat)
    k = _conv(sd, prefix + ".projection_k", feat)
    v = _conv(sd, prefix + ".projection_v", feat)
    msg = nonlocal_message(q, k, v, compat)
    if taps is not None:
        taps.append(dict(q=q, k=k, v=v, msg=msg))
    m = _conv(sd, prefix + ".fc_message.0", msg)
    m = F.relu(_bn(sd, prefix + ".fc_message.1", m))
    m = _conv(sd, prefix + ".fc_message.3", m)
    m = F.relu(_bn(sd, prefix + ".fc_message.4", m))
    m = _conv(sd, prefix + ".fc_message.6", m)
    return feat + m


# --------------------------------------------------------------------------
# R2  encoder  (models/PointDSC.py:65-77)
# --------------------------------------------------------------------------
def encoder(sd, corr_pos, compat, num_layers, taps=None):
    """corr_pos [1,N,6] -> feat [1,C,N]."""
    feat = _conv(sd, "encoder.layer0", corr_pos.permute(0, 2, 1))
    for i in range(num_layers):
        p = f"encoder.blocks.PointCN_layer_{i}"
        feat = F.relu(_bn(sd, p + ".1", _conv(sd, p + ".0", feat)))
        feat = nonlocal_block(sd, f"encoder.blocks.NonLocal_layer_{i}", feat, compat, taps)
    return feat


# --------------------------------------------------------------------------
# R4  confidence MLP  (models/PointDSC.py:107-113,171)
# --------------------------------------------------------------------------
def classification(sd, feat):
    """feat [1,C,N] -> raw logits [1,N]."""
    x = F.relu(_conv(sd, "classification.0", feat))
    x = F.relu(_conv(sd, "classification.2", x))
    return _conv(sd, "classification.4", x).squeeze(1)


# --------------------------------------------------------------------------
# R5  NMS seed picking  (models/PointDSC.py:199-217)
# --------------------------------------------------------------------------
def pick_seeds(src_dist, scores, R, max_num):
    assert scores.shape[0] == 1
    rel = scores.T >= scores
    rel = rel.bool() | (src_dist[0] >= R).bool()
    is_local_max = rel.min(-1)[0].float()
    keys = scores * is_local_max
    seeds = torch.argsort(keys, dim=1, descending=True)[:, 0:max_num]
    return seeds, is_local_max, keys


# --------------------------------------------------------------------------
# R6a  feature-space kNN  (models/common.py:48-69, ignore_self, normalized)
# --------------------------------------------------------------------------
def knn_all(normed, k):
    inner = 2 * torch.matmul(normed, normed.transpose(2, 1))
    dist = 2 - inner
    return dist.topk(k=k + 1, dim=-1, largest=False)[1][:, :, 1:], dist


# --------------------------------------------------------------------------
# R7  power iteration  (models/PointDSC.py:338-358)
# --------------------------------------------------------------------------
def leading_eigenvector(M, num_iterations):
    """M [S,k,k] -> v [S,k], iterations actually run (the allclose test spans
    the whole tensor, so every seed of the pair stops together)."""
    v = torch.ones_like(M[:, :, 0:1])
    last = v
    used = 0
    for _ in range(num_iterations):
        v = torch.bmm(M, v)
        v = v / (torch.norm(v, dim=1, keepdim=True) + 1e-6)
        used += 1
        if torch.allclose(v, last):
            break
        last = v
    return v.squeeze(-1), used


# --------------------------------------------------------------------------
# R8  weighted Kabsch  (models/common.py:7-45)
# --------------------------------------------------------------------------
def rigid_transform_3d(A, B, w):
    """A,B [bs,n,3], w [bs,n] -> T [bs,4,4].  torch.svd returns V (not V^T)."""
    bs = A.shape[0]
    w = w.clone()
    w[w < 0] = 0
    cA = torch.sum(A * w[:, :, None], dim=1, keepdim=True) / (torch.sum(w, dim=1, keepdim=True)[:, :, None] + 1e-6)
    cB = torch.sum(B * w[:, :, None], dim=1, keepdim=True) / (torch.sum(w, dim=1, keepdim=True)[:, :, None] + 1e-6)
    Am = A - cA
    Bm = B - cB
    H = Am.permute(0, 2, 1) @ torch.diag_embed(w) @ Bm
    U, S, V = torch.svd(H)
    d = torch.det(V @ U.permute(0, 2, 1))
    eye = torch.eye(3)[None].repeat(bs, 1, 1)
    eye[:, -1, -1] = d
    R = V @ eye @ U.permute(0, 2, 1)
    t = cB.permute(0, 2, 1) - R @ cA.permute(0, 2, 1)
    return integrate_trans(R, t), H


# --------------------------------------------------------------------------
# R6b + R7 + R8 + R9  per-seed hypotheses and scoring (models/PointDSC.py:234-336)
# --------------------------------------------------------------------------
def seed_hypotheses(seeds, normed, src, tgt, sigma, sigma_spat, k, num_iterations,
                    inlier_threshold):
    out = {}
    bs, N, C = normed.shape
    k = min(k, N - 1)
    knn_idx_all, _ = knn_all(normed, k)
    knn_idx = knn_idx_all.gather(1, seeds[:, :, None].expand(-1, -1, k))
    out["knn_idx"] = knn_idx
    flat = knn_idx.view(bs, -1)
    kf = normed.gather(1, flat[:, :, None].expand(-1, -1, C)).view(bs, -1, k, C)
    FM = torch.matmul(kf, kf.permute(0, 1, 3, 2))
    FM = torch.clamp(1 - (1 - FM) / sigma ** 2, min=0).view(-1, k, k)
    sk = src.gather(1, flat[:, :, None].expand(-1, -1, 3)).view(bs, -1, k, 3)
    tk = tgt.gather(1, flat[:, :, None].expand(-1, -1, 3)).view(bs, -1, k, 3)
    SM = ((sk[:, :, :, None, :] - sk[:, :, None, :, :]) ** 2).sum(-1) ** 0.5 \
        - ((tk[:, :, :, None, :] - tk[:, :, None, :, :]) ** 2).sum(-1) ** 0.5
    SM = torch.clamp(1 - SM ** 2 / sigma_spat ** 2, min=0).view(-1, k, k)
    M = FM * SM
    M[:, torch.arange(k), torch.arange(k)] = 0
    out["local_M"] = M
    v, used = leading_eigenvector(M, num_iterations)
    out["power_iters"] = used
    w = v.view(bs, -1, k)
    w = w / (torch.sum(w, dim=-1, keepdim=True) + 1e-6)
    w = w.view(-1, k)
    out["weights"] = w
    T, H = rigid_transform_3d(sk.view(-1, k, 3), tk.view(-1, k, 3), w)
    out["seed_H"] = H
    seed_trans = T.view(bs, -1, 4, 4)
    out["seed_trans"] = seed_trans
    pred = torch.einsum("bsnm,bmk->bsnk", seed_trans[:, :, :3, :3], src.permute(0, 2, 1)) \
        + seed_trans[:, :, :3, 3:4]
    pred = pred.permute(0, 1, 3, 2)
    L2 = torch.norm(pred - tgt[:, None, :, :], dim=-1)
    hit = L2 < inlier_threshold
    out["inlier_count"] = hit.sum(-1).to(torch.int32)
    fitness = torch.mean(hit.float(), dim=-1)
    out["seed_fitness"] = fitness
    best = fitness.argmax(dim=1)
    out["best_idx"] = best
    out["initial_trans"] = seed_trans.gather(1, best[:, None, None, None].expand(-1, -1, 4, 4)).squeeze(1)
    lab = L2.gather(1, best[:, None, None].expand(-1, -1, N)).squeeze(1)
    out["final_labels"] = (lab < inlier_threshold).float()
    return out


# --------------------------------------------------------------------------
# R10  post refinement  (models/PointDSC.py:403-438)
# --------------------------------------------------------------------------
def post_refinement(T, src, tgt, inlier_threshold):
    assert T.shape[0] == 1
    tau = 0.10 if inlier_threshold == 0.10 else 1.2   # quirk, lines 415-418
    prev = 0
    counts = []
    for _ in range(20):
        d = torch.norm(transform(src, T) - tgt, dim=-1)
        mask = (d < tau)[0]
        n = int(torch.sum(mask))
        counts.append(n)
        if abs(n - prev) < 1:
            break
        prev = n
        w = (1 / (1 + (d / tau) ** 2))[:, mask]
        T, _ = rigid_transform_3d(src[:, mask, :], tgt[:, mask, :], w)
    return T, counts


# --------------------------------------------------------------------------
# forward in test mode  (models/PointDSC.py:128-197)
# --------------------------------------------------------------------------
@torch.no_grad()
def forward_pair(sd, corr_pos, src, tgt, *, num_layers=12, num_iterations=10, ratio=0.1,
                 inlier_threshold=0.10, k=40, nms_radius=0.10, taps=False):
    """One pair ([1,N,.] tensors).  Returns every intermediate in a dict."""
    out = {}
    N = corr_pos.shape[1]
    sigma, sigma_spat = sd["sigma"], sd["sigma_spat"]
    src_dist, compat = spatial_compatibility(src, tgt, sigma_spat)
    layer_taps = [] if taps else None
    feat = encoder(sd, corr_pos, compat, num_layers, layer_taps).permute(0, 2, 1)
    normed = F.normalize(feat, p=2, dim=-1)
    conf = classification(sd, feat.permute(0, 2, 1))
    seeds, is_local_max, keys = pick_seeds(src_dist, conf, nms_radius, int(N * ratio))
    out.update(corr_features=feat, normed_features=normed, confidence=conf, seeds=seeds,
               is_local_max=is_local_max, seed_keys=keys)
    if taps:
        out["layer_taps"] = layer_taps
        out["compat"] = compat
    out.update(seed_hypotheses(seeds, normed, src, tgt, sigma, sigma_spat, k, num_iterations,
                               inlier_threshold))
    T, counts = post_refinement(out["initial_trans"], src, tgt, inlier_threshold)
    out["final_trans"] = T
    out["refine_counts"] = counts
    return out


@torch.no_grad()
def forward(sd, data, **cfg):
    """Batched convenience wrapper: B independent pairs.  Returns the
    reference's three keys (models/PointDSC.py:192-197) plus the per-pair dicts."""
    B = data["corr_pos"].shape[0]
    per = [forward_pair(sd, data["corr_pos"][b:b + 1], data["src_keypts"][b:b + 1],
                        data["tgt_keypts"][b:b + 1], **cfg) for b in range(B)]
    return {
        "final_trans": torch.cat([p["final_trans"] for p in per], 0),
        "final_labels": torch.cat([p["final_labels"] for p in per], 0),
        "M": None,
        "per_pair": per,
    }


# --------------------------------------------------------------------------
# R7b  full-matrix spectral matching  (cal_leading_eigenvector on the N x N compat with
# zero diagonal: the commented call at models/PointDSC.py:170 and
# baseline_scripts/baseline_3DMatch.py:19-46)
# --------------------------------------------------------------------------
@torch.no_grad()
def power_iteration_full(src, tgt, sigma_spat, num_iterations):
    _, compat = spatial_compatibility(src, tgt, sigma_spat)
    n = compat.shape[1]
    compat[:, torch.arange(n), torch.arange(n)] = 0
    return leading_eigenvector(compat, num_iterations)


# --------------------------------------------------------------------------
# metrics used by the parity tests  (libs/loss.py:44-49)
# --------------------------------------------------------------------------
def rotation_error_deg(Ra, Rb):
    """Angle between two rotations.  For rotations ||Ra-Rb||_F = 2*sqrt(2)*sin(angle/2); this
    form stays accurate near zero, where the trace/acos form of libs/loss.py:44-46 loses all
    digits in fp32 (acos(1-1e-7) is already 0.026 deg)."""
    d = torch.linalg.matrix_norm(Ra.double() - Rb.double())
    return torch.rad2deg(2 * torch.asin(torch.clamp(d / (2 * 2 ** 0.5), max=1.0)))


def translation_error(ta, tb):
    return torch.norm(ta.double() - tb.double(), dim=-1)
