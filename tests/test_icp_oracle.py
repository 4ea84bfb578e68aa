"""CPU known-answer checks of the ICP restatement (oracle/restate_icp.py).  open3d is absent: parity
with the reference's registration_icp call is unpinned (see the oracle's header)."""
import numpy as np

from oracle import restate_icp as I


def _rigid(rng, angle_deg, shift):
    ax = rng.normal(size=3); ax /= np.linalg.norm(ax)
    a = np.deg2rad(angle_deg)
    K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    T = np.eye(4)
    T[:3, :3] = np.eye(3) + np.sin(a) * K + (1 - np.cos(a)) * K @ K
    T[:3, 3] = shift
    return T


def test_umeyama_recovers_exact_motion():
    rng = np.random.default_rng(0)
    P = rng.normal(size=(50, 3))
    T = _rigid(rng, 37.0, [0.3, -0.2, 0.9])
    R, t = I.umeyama_rigid(P, P @ T[:3, :3].T + T[:3, 3])
    assert np.allclose(R, T[:3, :3], atol=1e-12) and np.allclose(t, T[:3, 3], atol=1e-12)
    R, t = I.umeyama_rigid(P[:0], P[:0])
    assert np.array_equal(R, np.eye(3)) and np.array_equal(t, np.zeros(3))


def test_icp_converges_to_the_exact_motion():
    rng = np.random.default_rng(1)
    src = rng.uniform(0, 3, (1500, 3))
    T = _rigid(rng, 25.0, [0.5, 0.1, -0.4])
    tgt = src @ T[:3, :3].T + T[:3, 3]
    T0 = _rigid(rng, 1.0, [0.01, -0.01, 0.005]) @ T
    Tr, fit, rmse, it, idx = I.icp_point_to_point(src, tgt, T0)
    assert fit == 1.0 and rmse < 1e-9 and (idx == np.arange(1500)).all()
    assert np.allclose(Tr, T, atol=1e-9) and 1 <= it <= 30


def test_icp_without_overlap_keeps_the_initial_transform():
    rng = np.random.default_rng(2)
    src = rng.uniform(0, 1, (100, 3))
    T0 = _rigid(rng, 10.0, [0.1, 0.2, 0.3])
    Tr, fit, rmse, it, idx = I.icp_point_to_point(src, src + 50.0, T0)
    assert fit == 0.0 and rmse == 0.0 and it == 1 and (idx == -1).all() and np.array_equal(Tr, T0)


def test_nearest_ties_go_to_the_lower_index():
    from scipy.spatial import cKDTree
    tgt = np.array([[1.0, 0, 0], [0.0, 0, 0], [0.0, 0, 0], [5.0, 5, 5]])
    idx, d2 = I.associate(np.array([[0.01, 0, 0], [9.0, 9, 9]]), tgt, cKDTree(tgt), 0.1)
    assert idx.tolist() == [1, -1] and np.isinf(d2[1])
