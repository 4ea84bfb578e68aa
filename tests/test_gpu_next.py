"""GPU parity of the callers either side of the hot path: correspondence construction and
evaluation metrics, through the C-ABI, against the reference-produced golden vectors and oracle/restate_next.py."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def cu(x):
    return x.cuda().contiguous()


def _check_nn(idx_got, idx_want, dist64, what, second_of=None):
    """Index equality except where the two candidates are closer than fp32 can tell apart (logged)."""
    got, want = idx_got.cpu().numpy(), np.asarray(idx_want)
    rows = np.arange(len(got))
    diff = np.nonzero(got != want)[0]
    gap = np.abs(dist64[rows, got] - dist64[rows, want])
    assert np.all(gap[diff] < 2e-6), f"{what}: {len(diff)} differences, worst gap {gap[diff].max():.3e}"
    if len(diff):
        print(f"{what}: {len(diff)} near-tie decisions differ (worst gap {gap[diff].max():.2e})")


@pytest.mark.parametrize("name", ["next_match_fcgf32", "next_match_fpfh33"])
def test_find_nn_and_mutual_vs_reference_golden(name):
    from pointdsc_b200 import matching
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    F0, F1 = torch.from_numpy(g["F0"]), torch.from_numpy(g["F1"])
    d64 = torch.cdist(F0.double(), F1.double()).numpy()
    i0, i1, i2 = matching.find_nn(cu(F0), cu(F1), return_2nd=True)
    assert torch.equal(i0.cpu(), torch.arange(F0.shape[0]))
    _check_nn(i1, g["idx1"], d64, "nearest")
    _check_nn(i2, g["idx2"], d64, "second nearest")
    assert bool((i1 != i2).all())
    m0, m1 = matching.nn_to_mutual(cu(F0), cu(F1), i0, i1)
    got = set(zip(m0.cpu().tolist(), m1.cpu().tolist()))
    want = set(zip(g["mutual0"].tolist(), g["mutual1"].tolist()))
    assert got == want, f"mutual sets differ by {len(got ^ want)} pairs"
    assert bool((m0[1:] > m0[:-1]).all())                      # ascending source index, like the reference
    m0b, m1b, sec = matching.nn_to_mutual(cu(F0), cu(F1), i0, i1, idx1_2nd=i2)
    assert torch.equal(sec, i2[m0b])


@pytest.mark.parametrize("use_mutual", [True, False])
@pytest.mark.parametrize("shape", [(700, 650, 32), (37, 5, 33), (130, 1000, 7), (1, 3, 64)])
def test_lidar_mutual_correspondences_vs_oracle(shape, use_mutual):
    from pointdsc_b200 import matching
    from oracle import restate_next
    from oracle.gen_golden_next import descriptors
    N0, N1, D = shape
    F0, F1 = descriptors(100 + N0, N0, N1, D)
    want, dist = restate_next.lidar_mutual_corr(F0.numpy(), F1.numpy(), use_mutual=use_mutual)
    got = matching.mutual_nn_correspondences(cu(F0), cu(F1), use_mutual=use_mutual).cpu().numpy()
    if got.shape == want.shape and np.array_equal(got, want):
        return
    # any difference must come from a decision closer than fp32 resolves
    d64 = np.sqrt(np.maximum(2 - 2 * (F0.double() @ F1.double().T).numpy() + 1e-6, 0))
    s = np.sort(d64, axis=1)
    near_rows = set(np.nonzero(s[:, min(1, N1 - 1)] - s[:, 0] < 2e-6)[0].tolist())
    sc = np.sort(d64, axis=0)
    near_cols = set(np.nonzero(sc[min(1, N0 - 1)] - sc[0] < 2e-6)[0].tolist())
    for i, j in set(map(tuple, got.tolist())) ^ set(map(tuple, want.tolist())):
        assert i in near_rows or j in near_cols, f"pair ({i},{j}) differs without a near tie"


def test_nn_match_large_batched_properties():
    """Sizes of the LiDAR front end (12 000 x 32-dim FCGF), two problems per call: every answer is a true minimum of
    a float64 recomputation (up to fp32 resolution), and the mutual set is symmetric under swapping the clouds."""
    from pointdsc_b200 import _cabi
    from oracle.gen_golden_next import descriptors
    P, N0, N1, D = 2, 12000, 11000, 32
    Fs = [descriptors(7 + p, N0, N1, D) for p in range(P)]
    A = cu(torch.stack([f[0] for f in Fs]))
    B = cu(torch.stack([f[1] for f in Fs]))
    iab, _, v1, v2 = _cabi.nn_match(A, B, metric=_cabi.MATCH_DOT, want_values=True)
    iba, _, _, _ = _cabi.nn_match(B, A, metric=_cabi.MATCH_DOT)
    torch.cuda.synchronize()
    assert bool((v1 <= v2).all())
    for p in range(P):
        d = torch.sqrt(torch.clamp(2 - 2 * (A[p].double() @ B[p].double().T) + 1e-6, min=0))
        best = d.min(dim=1).values
        got = d[torch.arange(N0, device="cuda"), iab[p]]
        assert float((got - best).max()) < 2e-6
        assert float((v1[p].double() - best).abs().max()) < 2e-6
    cab, nab = _cabi.mutual_corr(iab, iba, N1)
    cba, nba = _cabi.mutual_corr(iba, iab, N0)
    for p in range(P):
        ab = set(map(tuple, cab[p, : int(nab[p])].cpu().tolist()))
        ba = set((j, i) for i, j in cba[p, : int(nba[p])].cpu().tolist())
        assert ab == ba and len(ab) > 1000


def test_registration_metrics_vs_reference_golden_and_oracle():
    from pointdsc_b200 import metrics
    from oracle import restate_next
    g = np.load(os.path.join(GOLDEN, "next_metrics.npz"))
    t = lambda k: torch.from_numpy(g[k])
    got = metrics.registration_errors(cu(t("est")), cu(t("gt")), cu(t("src")), cu(t("tgt"))).cpu().double()
    want = restate_next.transformation_metrics(t("est"), t("gt"), t("src"), t("tgt"))
    ref = g["trans_loss"]
    # RE: acos of a fp32 cosine near 1; d(angle) = d(cos)/sin(angle), with |d cos| <= 3e-7
    tol_re = np.degrees(3e-7 / np.sin(np.radians(want[:, 0].numpy()))) + 1e-4
    assert np.all(np.abs(got[:, 0].numpy() - want[:, 0].numpy()) < tol_re)
    assert np.all(np.abs(got[:, 0].numpy() - ref[:, 2]) < 2 * tol_re)
    assert torch.allclose(got[:, 1:], want[:, 1:], rtol=1e-5, atol=1e-5)
    assert np.allclose(got[:, 1].numpy(), ref[:, 3], rtol=1e-5, atol=1e-4)
    assert np.allclose(got[:, 2].numpy(), ref[:, 4], rtol=1e-5)
    assert np.allclose(got[:, 3].numpy(), ref[:, 0], rtol=1e-5)
    # the module form, batch of one pair, against the reference's own return values
    tl = metrics.TransformationLoss(re_thre=15, te_thre=30)
    for b in range(3):
        loss, recall, RE, TE, RMSE = tl(cu(t("est")[b:b + 1]), cu(t("gt")[b:b + 1]), cu(t("src")[b:b + 1]), cu(t("tgt")[b:b + 1]),
                                        cu(t("logits")[b:b + 1]))
        assert abs(float(loss) - ref[b, 0]) < 1e-5 * max(1.0, ref[b, 0]) and float(recall) == ref[b, 1]
        assert abs(float(TE) - ref[b, 3]) < 1e-4 and abs(float(RMSE) - ref[b, 4]) < 1e-6
    # a pair with no predicted inlier contributes zero loss (libs/loss.py:57-59)
    loss0 = tl(cu(t("est")[:1]), cu(t("gt")[:1]), cu(t("src")[:1]), cu(t("tgt")[:1]), cu(-torch.ones(1, 400)))[0]
    assert float(loss0) == 0.0


@pytest.mark.parametrize("N", [1, 31, 400, 5000])
def test_classification_stats_vs_oracle(N):
    from pointdsc_b200 import metrics
    from oracle import restate_next
    g = torch.Generator().manual_seed(N)
    B = 4
    gt = (torch.rand(B, N, generator=g) < 0.3).float()
    pred = torch.randn(B, N, generator=g) + 1.2 * (gt - 0.5)
    pred[0] = -1.0                       # no predicted inlier: precision 0 by sklearn's convention
    gt[1] = 0.0                          # no true inlier: recall 0
    s = metrics.classification_stats(cu(pred), cu(gt))
    for b in range(B):
        w = restate_next.classification_stats(pred[b], gt[b])
        assert s["counts"][b].cpu().tolist() == [w["tp"], w["fp"], w["fn"], w["tn"]]
        for k in ("precision", "recall", "f1"):
            assert abs(float(s[k][b]) - w[k]) < 1e-12
        for k in ("logit_true", "logit_false"):
            assert abs(float(s[k][b]) - w[k]) < 1e-5 * max(1.0, abs(w[k]))


def test_golden_classification_stats():
    from pointdsc_b200 import metrics
    g = np.load(os.path.join(GOLDEN, "next_metrics.npz"))
    s = metrics.classification_stats(cu(torch.from_numpy(g["logits"])), cu(torch.from_numpy(g["labels"])))
    want = g["cls_stats"]                # sklearn precision / recall / f1 + mean logits from the reference
    for b in range(3):
        assert abs(float(s["precision"][b]) - want[b, 0]) < 1e-12
        assert abs(float(s["recall"][b]) - want[b, 1]) < 1e-12
        assert abs(float(s["f1"][b]) - want[b, 2]) < 1e-12
        assert abs(float(s["logit_true"][b]) - want[b, 3]) < 1e-5
        assert abs(float(s["logit_false"][b]) - want[b, 4]) < 1e-5
