"""CPU checks of the front-end restatement (oracle/restate_frontend.py): known answers, the committed
crop of the reference's demo clouds, and the host-side PLY reader.  open3d is absent, so these are known-answer and
regression tests, not comparisons with the reference's open3d calls (parity unpinned, see the oracle's header)."""
import os
import struct

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from oracle import restate_frontend as F


def _golden():
    return np.load(os.path.join(GOLDEN_DIR, "front_demo_crop.npz"))


def test_hybrid_search_matches_brute_force():
    rng = np.random.default_rng(0)
    pts = rng.uniform(0, 1, (400, 3))
    pts[50] = pts[10]                                   # an exact duplicate: a zero-distance tie
    idx, d2, cnt = F.hybrid_search(pts, 0.2, 12)
    D = ((pts[:, None] - pts[None]) ** 2).sum(-1)
    for i in (0, 10, 50, 123, 399):
        order = np.lexsort((np.arange(400), D[i]))
        order = order[D[i][order] < 0.04][:12]
        assert cnt[i] == len(order)
        assert (idx[i, :cnt[i]] == order).all()
        assert np.allclose(d2[i, :cnt[i]], D[i][order], rtol=1e-12)
        assert (idx[i, cnt[i]:] == -1).all() and np.isinf(d2[i, cnt[i]:]).all()
    assert idx[10, 0] == 10 and idx[10, 1] == 50 and idx[50, 0] == 10      # ties resolve to the lower index


def test_smallest_eigenvector_matches_lapack():
    rng = np.random.default_rng(1)
    A = rng.normal(size=(2000, 3, 3))
    C = A @ A.transpose(0, 2, 1) * rng.uniform(1e-6, 1e3, (2000, 1, 1))
    C[:50] = np.diag([3.0, 1.0, 2.0])                   # diagonal branch
    C[50:60] = 0.0                                       # zero matrix -> zero vector
    v = F.smallest_eigenvector(C)
    w, V = np.linalg.eigh(C[60:])
    gap = (w[:, 1] - w[:, 0]) / w[:, 2]
    ok = gap > 1e-6
    dots = np.abs((v[60:] * V[:, :, 0]).sum(1))
    assert np.all(dots[ok] > 1 - 1e-9)
    assert np.allclose(np.linalg.norm(v[60:], axis=1), 1.0, atol=1e-12)
    assert (v[:50] == np.array([0.0, 1.0, 0.0])).all() and (v[50:60] == 0).all()


def test_normals_of_a_noisy_plane():
    rng = np.random.default_rng(2)
    n = np.array([1.0, 2.0, -2.0]) / 3.0
    u = np.cross(n, [0, 0, 1.0]); u /= np.linalg.norm(u)
    v = np.cross(n, u)
    ab = rng.uniform(-1, 1, (3000, 2))
    pts = ab[:, :1] * u + ab[:, 1:] * v + rng.normal(0, 1e-3, (3000, 1)) * n
    idx, d2, cnt = F.hybrid_search(pts, 0.15, 30)
    nrm = F.estimate_normals(pts, idx, cnt)
    assert np.all(np.abs(nrm @ n) > 0.995)
    lonely = np.vstack([pts, [[50.0, 50.0, 50.0]]])     # fewer than 3 neighbours -> (0,0,1)
    idx, d2, cnt = F.hybrid_search(lonely, 0.15, 30)
    assert cnt[-1] == 1 and (F.estimate_normals(lonely, idx, cnt)[-1] == [0, 0, 1]).all()


def test_voxel_down_sample_known_answer():
    pts = np.array([[0.0, 0.0, 0.0], [0.1, 0.1, 0.1], [1.0, 0.0, 0.0], [0.9, 0.1, 0.0], [0.0, 1.0, 5.0]])
    nrm = np.array([[0, 0, 1.0], [0, 0, 1.0], [1.0, 0, 0], [0, 1.0, 0], [0, 0, -2.0]])
    p, n = F.voxel_down_sample(pts, nrm, 0.5)
    # voxel origin = min - 0.25; ascending (ix,iy,iz)
    assert np.allclose(p, [[0.05, 0.05, 0.05], [0.0, 1.0, 5.0], [0.95, 0.05, 0.0]])
    assert np.allclose(n, [[0, 0, 1], [0, 0, -1], [2 ** -0.5, 2 ** -0.5, 0]])


def test_fpfh_histogram_mass():
    g = _golden()
    pts, nrm = g["src_down_pts"], g["src_down_normals"]
    idx, d2, cnt = F.hybrid_search(pts, 0.25, 100)
    assert (cnt == g["src_down_cnt"]).all()
    f = F.compute_fpfh(pts, nrm, idx, d2, cnt)
    full = cnt > 1
    for grp in range(3):
        assert np.allclose(f[full, 11 * grp:11 * grp + 11].sum(1), 200.0, atol=1e-9)   # 100 (SPFH) + 100 (weighted)
    assert np.allclose(f, g["src_fpfh"], atol=2e-4)                                      # stored as float32


def test_oracle_regenerates_demo_crop():
    g = _golden()
    pts = g["tgt_raw"].astype(np.float64)[:6000]
    idx, d2, cnt = F.hybrid_search(pts, 0.1, 30)
    assert cnt.max() <= 30 and (np.diff(d2[:, :5], axis=1) >= 0).all()
    # the first 6000 points are a subset of the crop, so only self-consistency is checked here; the full crop below
    pts = g["src_raw"].astype(np.float64)
    sel = g["src_sel"][:200]
    sub_idx, sub_d2, sub_cnt = F.hybrid_search(pts, 0.1, 30)
    assert (sub_idx[sel] == g["src_sel_idx"][:200]).all()
    nrm = F.estimate_normals(pts, sub_idx, sub_cnt)
    assert np.allclose(nrm[g["src_sel"]], g["src_sel_normals"], atol=1e-12)
    dp, dn = F.voxel_down_sample(pts, nrm, 0.05)
    assert np.array_equal(dp, g["src_down_pts"]) and np.allclose(dn, g["src_down_normals"], atol=1e-12)


def test_matching_lines():
    rng = np.random.default_rng(3)
    a = rng.normal(size=(60, 33)); a /= np.linalg.norm(a, axis=1, keepdims=True)
    b = rng.normal(size=(45, 33)); b /= np.linalg.norm(b, axis=1, keepdims=True)
    b[7] = b[3]                                          # equal descriptors: the first index wins
    a[5] = b[3]
    ap, bp = rng.normal(size=(60, 3)), rng.normal(size=(45, 3))
    corr, sk, tk, cp = F.match_and_centre(ap, a, bp, b)
    assert corr[5, 1] == 3 and (corr[:, 0] == np.arange(60)).all()
    assert (corr[:, 1] == np.argmin(np.sqrt(2 - 2 * (a @ b.T) + 1e-6), axis=1)).mean() > 0.95
    assert np.allclose(cp.mean(0), 0, atol=1e-12) and np.allclose(cp - cp[0], np.hstack([sk, tk]) - np.hstack([sk, tk])[0])


def _write_ply(path, xyz, fmt, extra=False):
    n = len(xyz)
    props = "property float x\nproperty float y\nproperty float z\n"
    if extra:
        props = "property uchar red\n" + props + "property double w\n"
    head = f"ply\nformat {fmt} 1.0\ncomment test\nelement vertex {n}\n{props}element face 0\nproperty list uchar int vertex_indices\nend_header\n"
    with open(path, "wb") as f:
        f.write(head.encode())
        if fmt == "ascii":
            for p in xyz:
                row = ([7] if extra else []) + [repr(float(v)) for v in p] + ([0.5] if extra else [])
                f.write((" ".join(str(v) for v in row) + "\n").encode())
        else:
            e = "<" if fmt == "binary_little_endian" else ">"
            for p in xyz:
                if extra:
                    f.write(struct.pack(e + "B", 7))
                f.write(struct.pack(e + "fff", *p))
                if extra:
                    f.write(struct.pack(e + "d", 0.5))


@pytest.mark.parametrize("fmt", ["ascii", "binary_little_endian", "binary_big_endian"])
@pytest.mark.parametrize("extra", [False, True])
def test_ply_readers(tmp_path, fmt, extra):
    from pointdsc_b200 import frontend
    xyz = np.random.default_rng(4).normal(size=(17, 3)).astype(np.float32)
    path = str(tmp_path / "c.ply")
    _write_ply(path, xyz, fmt, extra)
    assert np.array_equal(F.read_ply_xyz(path), xyz.astype(np.float64))
    assert np.array_equal(frontend.read_point_cloud(path, device="cpu").numpy(), xyz.astype(np.float64))
    with open(path, "wb") as f:
        f.write(b"not a ply\n")
    with pytest.raises(ValueError):
        frontend.read_point_cloud(path, device="cpu")
