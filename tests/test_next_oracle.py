"""CPU: the restatement of the callers either side of the hot path (oracle/restate_next.py) against the golden
vectors produced by the unmodified reference (oracle/gen_golden_next.py)."""
import os

import numpy as np
import pytest
import torch

from oracle import restate_next

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("name", ["next_match_fcgf32", "next_match_fpfh33"])
def test_find_nn_and_mutual_match_reference(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    F0, F1 = torch.from_numpy(g["F0"]), torch.from_numpy(g["F1"])
    i0, i1, i2, _ = restate_next.find_nn(F0, F1, return_2nd=True)
    assert np.array_equal(i1.numpy(), g["idx1"])
    assert np.array_equal(i2.numpy(), g["idx2"])
    m0, m1 = restate_next.nn_to_mutual(F0, F1, i0, i1)
    assert np.array_equal(m0.numpy(), g["mutual0"]) and np.array_equal(m1.numpy(), g["mutual1"])


def test_lidar_mutual_is_the_mutual_subset():
    g = np.load(os.path.join(GOLDEN, "next_match_fcgf32.npz"))
    corr, dist = restate_next.lidar_mutual_corr(g["F0"], g["F1"], use_mutual=True)
    allc, _ = restate_next.lidar_mutual_corr(g["F0"], g["F1"], use_mutual=False)
    assert allc.shape == (g["F0"].shape[0], 2)
    assert np.all(np.diff(corr[:, 0]) > 0)
    for i, j in corr:
        assert dist[i, j] == dist[i].min() and dist[i, j] == dist[:, j].min()
    # on normalised descriptors the two distance forms agree on every decision that is not a near tie
    same = allc[:, 1] == g["idx1"]
    gap = np.sort(dist, axis=1)
    assert np.all(same | (gap[:, 1] - gap[:, 0] < 1e-5))


def test_metrics_match_reference():
    g = np.load(os.path.join(GOLDEN, "next_metrics.npz"))
    t = lambda k: torch.from_numpy(g[k])
    m = restate_next.transformation_metrics(t("est"), t("gt"), t("src"), t("tgt"))
    ref = g["trans_loss"]                      # loss, recall, RE, TE, RMSE per pair (fp32 reference)
    assert np.allclose(m[:, 0].numpy(), ref[:, 2], rtol=0, atol=2e-3)      # RE [deg]: fp32 acos near 1
    assert np.allclose(m[:, 1].numpy(), ref[:, 3], rtol=1e-5, atol=1e-4)   # TE [cm]
    assert np.allclose(m[:, 2].numpy(), ref[:, 4], rtol=1e-5)              # RMSE
    assert np.allclose(m[:, 3].numpy(), ref[:, 0], rtol=1e-5)              # loss = mean squared residual
    assert np.all(ref[:, 1] == 100.0)
    for b in range(3):
        s = restate_next.classification_stats(t("logits")[b], t("labels")[b])
        want = g["cls_stats"][b]
        assert abs(s["precision"] - want[0]) < 1e-12 and abs(s["recall"] - want[1]) < 1e-12 and abs(s["f1"] - want[2]) < 1e-12
        assert abs(s["logit_true"] - want[3]) < 1e-5 and abs(s["logit_false"] - want[4]) < 1e-5
