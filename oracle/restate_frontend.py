"""CPU restatement of the demo front end: PLY -> normals -> voxel down-sample -> FPFH-33 ->
nearest-neighbour matching -> centred corr_pos.  TEST INFRASTRUCTURE (see oracle/__init__.py): imported by tests/
and scripts that time the CPU leg only.

**Parity unpinned.**  The reference does these steps through open3d (`demo_registration.py:37-44`,
`utils/pointcloud.py:20-21`; pinned `open3d=0.9.0.0`, `environment.yml:76`), a third-party dependency that is not
vendored in /root/reference and is not installed here, and the reference holds no golden vectors for it.  What
follows restates the published algorithms open3d implements, in float64 as open3d does:

  hybrid search      FLANN radius search capped at max_nn: the max_nn nearest points with dist^2 < radius^2, the
                     query itself included, ascending distance.  Ties: ascending index (FLANN leaves them open).
  estimate_normals   covariance from the nine running sums (E[xx] - E[x]E[x]), eigenvector of the smallest eigenvalue
                     by the non-iterative solver of D. Eberly, "A Robust Eigensolver for 3x3 Symmetric Matrices"
                     (the sign of the result is whatever the cross products give: open3d does not orient normals).
                     Fewer than 3 neighbours -> (0,0,1); a zero eigenvector -> (0,0,1).
  voxel_down_sample  voxel index floor((p - (min_bound - voxel/2)) / voxel); mean point and normalised summed normal
                     per occupied voxel.  open3d emits the voxels in hash-map order; here ascending (ix, iy, iz).
  FPFH               R. Rusu et al., "Fast Point Feature Histograms (FPFH) for 3D Registration", ICRA 2009, in the
                     PCL/open3d form: SPFH = three 11-bin histograms of the Darboux-frame angles over the neighbours
                     (increment 100/(m-1)); FPFH = SPFH + the 1/dist^2-weighted neighbour SPFHs, each 11-bin group
                     rescaled to sum 100.
  matching           demo_registration.py:101-108, exactly as written there (float64).

The arithmetic order (sums over neighbours in ascending (dist^2, index) order, dist^2 = dx*dx + dy*dy + dz*dz without
fused multiply-add) is part of the restatement: the CUDA kernels follow the same order, so neighbour lists are compared
bit-exactly and floating-point outputs to ~1e-12.
"""
from __future__ import annotations

import numpy as np


# ----------------------------------------------------------------------------------------------- PLY
def read_ply_xyz(path):
    """Vertex positions of a PLY file (binary little/big endian or ascii; any extra per-vertex scalar properties
    are skipped) as float64 [n,3] -- what o3d.io.read_point_cloud(...).points holds (demo_registration.py:38)."""
    np_types = {"char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2",
                "ushort": "u2", "uint16": "u2", "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4",
                "float": "f4", "float32": "f4", "double": "f8", "float64": "f8"}
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError("not a PLY file")
        fmt, n, props, in_vertex = None, None, [], False
        while True:
            line = f.readline()
            if not line:
                raise ValueError("PLY header not terminated")
            tok = line.decode("ascii", "replace").split()
            if not tok:
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                in_vertex = tok[1] == "vertex"
                if in_vertex:
                    n = int(tok[2])
            elif tok[0] == "property" and in_vertex:
                if tok[1] == "list":
                    raise ValueError("list property on vertices is not supported")
                props.append((tok[2], np_types[tok[1]]))
            elif tok[0] == "end_header":
                break
        if fmt is None or n is None:
            raise ValueError("PLY header has no format / vertex element")
        names = [p[0] for p in props]
        if fmt == "ascii":
            rows = np.loadtxt(f, max_rows=n, ndmin=2)
            return np.stack([rows[:, names.index(c)] for c in "xyz"], 1).astype(np.float64)
        end = "<" if fmt == "binary_little_endian" else ">"
        dt = np.dtype([(nm, end + ty) for nm, ty in props])
        rec = np.frombuffer(f.read(n * dt.itemsize), dtype=dt, count=n)
        return np.stack([rec[c].astype(np.float64) for c in "xyz"], 1)


# ----------------------------------------------------------------------------------------------- search
def _dist2(q, p):
    d = p - q
    return d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1] + d[..., 2] * d[..., 2]


def hybrid_search(points, radius, max_nn):
    """idx [n,max_nn] int32 (-1 padded), dist2 [n,max_nn] float64 (inf padded), count [n] int32."""
    from scipy.spatial import cKDTree
    points = np.ascontiguousarray(points, dtype=np.float64)
    n = len(points)
    idx = np.full((n, max_nn), -1, np.int32)
    d2o = np.full((n, max_nn), np.inf, np.float64)
    cnt = np.zeros(n, np.int32)
    tree = cKDTree(points)
    r2 = radius * radius
    cand = tree.query_ball_point(points, radius * (1 + 1e-9) + 1e-12, workers=-1)
    for i in range(n):
        c = np.asarray(cand[i], dtype=np.int64)
        d2 = _dist2(points[i], points[c])
        keep = d2 < r2
        c, d2 = c[keep], d2[keep]
        order = np.lexsort((c, d2))[:max_nn]
        m = len(order)
        idx[i, :m] = c[order]
        d2o[i, :m] = d2[order]
        cnt[i] = m
    return idx, d2o, cnt


# ----------------------------------------------------------------------------------------------- normals
def _cross(a, b):
    return np.stack([a[:, 1] * b[:, 2] - a[:, 2] * b[:, 1], a[:, 2] * b[:, 0] - a[:, 0] * b[:, 2],
                     a[:, 0] * b[:, 1] - a[:, 1] * b[:, 0]], 1)


def _dot(a, b):
    return a[:, 0] * b[:, 0] + a[:, 1] * b[:, 1] + a[:, 2] * b[:, 2]


def _evec0(A, ev):
    """Eberly, eigenvector of A for the (well separated) eigenvalue ev: the largest cross product of two rows of A - ev I."""
    r0 = np.stack([A[:, 0, 0] - ev, A[:, 0, 1], A[:, 0, 2]], 1)
    r1 = np.stack([A[:, 0, 1], A[:, 1, 1] - ev, A[:, 1, 2]], 1)
    r2 = np.stack([A[:, 0, 2], A[:, 1, 2], A[:, 2, 2] - ev], 1)
    c01, c02, c12 = _cross(r0, r1), _cross(r0, r2), _cross(r1, r2)
    d0, d1, d2 = _dot(c01, c01), _dot(c02, c02), _dot(c12, c12)
    out = c01 / np.sqrt(d0)[:, None]
    dmax = d0.copy()
    sel = d1 > dmax
    out[sel] = (c02 / np.sqrt(d1)[:, None])[sel]
    dmax = np.where(sel, d1, dmax)
    sel = d2 > dmax
    out[sel] = (c12 / np.sqrt(d2)[:, None])[sel]
    return out


def _evec1(A, w, ev):
    """Eberly, eigenvector for the second eigenvalue, searched in the plane orthogonal to w."""
    big0 = np.abs(w[:, 0]) > np.abs(w[:, 1])
    inv_a = 1.0 / np.sqrt(w[:, 0] * w[:, 0] + w[:, 2] * w[:, 2])
    inv_b = 1.0 / np.sqrt(w[:, 1] * w[:, 1] + w[:, 2] * w[:, 2])
    z = np.zeros_like(inv_a)
    U = np.where(big0[:, None], np.stack([-w[:, 2] * inv_a, z, w[:, 0] * inv_a], 1),
                 np.stack([z, w[:, 2] * inv_b, -w[:, 1] * inv_b], 1))
    V = _cross(w, U)

    def mv(x):
        return np.stack([A[:, 0, 0] * x[:, 0] + A[:, 0, 1] * x[:, 1] + A[:, 0, 2] * x[:, 2],
                         A[:, 0, 1] * x[:, 0] + A[:, 1, 1] * x[:, 1] + A[:, 1, 2] * x[:, 2],
                         A[:, 0, 2] * x[:, 0] + A[:, 1, 2] * x[:, 1] + A[:, 2, 2] * x[:, 2]], 1)
    AU, AV = mv(U), mv(V)
    m00, m01, m11 = _dot(U, AU) - ev, _dot(U, AV), _dot(V, AV) - ev
    a00, a01, a11 = np.abs(m00), np.abs(m01), np.abs(m11)
    out = U.copy()
    with np.errstate(divide="ignore", invalid="ignore"):
        # branch 1: a00 >= a11
        t = m01 / m00; s = 1.0 / np.sqrt(1 + t * t); r1a = (t * s)[:, None] * U - s[:, None] * V        # a00 >= a01
        t = m00 / m01; s = 1.0 / np.sqrt(1 + t * t); r1b = s[:, None] * U - (t * s)[:, None] * V        # a00 <  a01
        # branch 2: a00 < a11
        t = m01 / m11; s = 1.0 / np.sqrt(1 + t * t); r2a = s[:, None] * U - (t * s)[:, None] * V        # a11 >= a01
        t = m11 / m01; s = 1.0 / np.sqrt(1 + t * t); r2b = (t * s)[:, None] * U - s[:, None] * V        # a11 <  a01
    b1 = a00 >= a11
    nz1 = np.maximum(a00, a01) > 0
    nz2 = np.maximum(a11, a01) > 0
    sel = b1 & nz1 & (a00 >= a01); out[sel] = r1a[sel]
    sel = b1 & nz1 & (a00 < a01); out[sel] = r1b[sel]
    sel = ~b1 & nz2 & (a11 >= a01); out[sel] = r2a[sel]
    sel = ~b1 & nz2 & (a11 < a01); out[sel] = r2b[sel]
    return out


def smallest_eigenvector(C):
    """Unit eigenvector of the smallest eigenvalue of symmetric C [n,3,3] (Eberly's non-iterative solver)."""
    C = np.asarray(C, np.float64)
    n = len(C)
    out = np.zeros((n, 3))
    mx = C.reshape(n, 9).max(1)
    ok = mx != 0
    A = C / np.where(ok, mx, 1.0)[:, None, None]
    norm = A[:, 0, 1] ** 2 + A[:, 0, 2] ** 2 + A[:, 1, 2] ** 2
    diag = ok & ~(norm > 0)
    gen = ok & (norm > 0)
    # diagonal matrices: the axis of the strictly smallest entry, z otherwise
    a, b, c = A[:, 0, 0], A[:, 1, 1], A[:, 2, 2]
    dsel = np.where((a < b) & (a < c), 0, np.where((b < a) & (b < c), 1, 2))
    out[diag, dsel[diag]] = 1.0
    if gen.any():
        G = A[gen]
        q = (G[:, 0, 0] + G[:, 1, 1] + G[:, 2, 2]) / 3.0
        b00, b11, b22 = G[:, 0, 0] - q, G[:, 1, 1] - q, G[:, 2, 2] - q
        p = np.sqrt((b00 * b00 + b11 * b11 + b22 * b22 + norm[gen] * 2.0) / 6.0)
        c00 = b11 * b22 - G[:, 1, 2] * G[:, 1, 2]
        c01 = G[:, 0, 1] * b22 - G[:, 1, 2] * G[:, 0, 2]
        c02 = G[:, 0, 1] * G[:, 1, 2] - b11 * G[:, 0, 2]
        det = (b00 * c00 - G[:, 0, 1] * c01 + G[:, 0, 2] * c02) / (p * p * p)
        half = np.clip(det * 0.5, -1.0, 1.0)
        ang = np.arccos(half) / 3.0
        beta2 = np.cos(ang) * 2.0
        beta0 = np.cos(ang + 2.09439510239319549) * 2.0
        beta1 = -(beta0 + beta2)
        e0, e1, e2 = q + p * beta0, q + p * beta1, q + p * beta2
        pos = half >= 0
        # half >= 0: e2 is the isolated one; else e0
        first = _evec0(G, np.where(pos, e2, e0))
        second = _evec1(G, first, e1)
        third = np.where(pos[:, None], _cross(second, first), _cross(first, second))
        efirst = np.where(pos, e2, e0)
        eother = np.where(pos, e0, e2)
        res = third
        sel1 = (e1 < efirst) & (e1 < eother)
        res = np.where(sel1[:, None], second, res)
        sel0 = (efirst < e1) & (efirst < eother)
        res = np.where(sel0[:, None], first, res)
        out[gen] = res
    return out


def estimate_normals(points, idx, cnt):
    """open3d EstimateNormals on hybrid-search results (utils/pointcloud.py:20-21): float64 [n,3]."""
    points = np.asarray(points, np.float64)
    n, max_nn = idx.shape
    cum = np.zeros((n, 9))
    for k in range(max_nn):
        valid = k < cnt
        p = points[np.where(valid, idx[:, k], 0)]
        terms = np.stack([p[:, 0], p[:, 1], p[:, 2], p[:, 0] * p[:, 0], p[:, 0] * p[:, 1], p[:, 0] * p[:, 2],
                          p[:, 1] * p[:, 1], p[:, 1] * p[:, 2], p[:, 2] * p[:, 2]], 1)
        cum += np.where(valid[:, None], terms, 0.0)
    cum /= np.maximum(cnt, 1)[:, None].astype(np.float64)
    C = np.empty((n, 3, 3))
    C[:, 0, 0] = cum[:, 3] - cum[:, 0] * cum[:, 0]
    C[:, 1, 1] = cum[:, 6] - cum[:, 1] * cum[:, 1]
    C[:, 2, 2] = cum[:, 8] - cum[:, 2] * cum[:, 2]
    C[:, 0, 1] = C[:, 1, 0] = cum[:, 4] - cum[:, 0] * cum[:, 1]
    C[:, 0, 2] = C[:, 2, 0] = cum[:, 5] - cum[:, 0] * cum[:, 2]
    C[:, 1, 2] = C[:, 2, 1] = cum[:, 7] - cum[:, 1] * cum[:, 2]
    nrm = smallest_eigenvector(C)
    bad = (cnt < 3) | ((nrm * nrm).sum(1) == 0)
    nrm[bad] = np.array([0.0, 0.0, 1.0])
    return nrm


# ----------------------------------------------------------------------------------------------- voxels
def voxel_down_sample(points, normals, voxel):
    """Mean point and normalised summed normal per occupied voxel, voxels in ascending (ix,iy,iz) order, points summed
    in ascending input index (as open3d's loop does)."""
    points = np.asarray(points, np.float64)
    origin = points.min(0) - voxel * 0.5
    vi = np.floor((points - origin) / voxel).astype(np.int64)
    dims = vi.max(0) + 1
    key = (vi[:, 0] * dims[1] + vi[:, 1]) * dims[2] + vi[:, 2]
    order = np.argsort(key, kind="stable")
    uk, start, count = np.unique(key[order], return_index=True, return_counts=True)
    m = len(uk)
    ps = np.zeros((m, 3))
    ns = np.zeros((m, 3))
    for j in range(int(count.max())):
        live = j < count
        src = order[np.where(live, start + j, 0)]
        ps += np.where(live[:, None], points[src], 0.0)
        if normals is not None:
            ns += np.where(live[:, None], normals[src], 0.0)
    ps /= count[:, None].astype(np.float64)
    if normals is None:
        return ps, None
    z = np.sqrt(ns[:, 0] * ns[:, 0] + ns[:, 1] * ns[:, 1] + ns[:, 2] * ns[:, 2])
    ns = np.where((z > 0)[:, None], ns / np.where(z > 0, z, 1.0)[:, None], ns)
    return ps, ns


# ----------------------------------------------------------------------------------------------- FPFH
def _pair_features(p1, n1, p2, n2):
    """Darboux-frame angles (theta, alpha, phi) of a point pair and the validity mask."""
    dp = p2 - p1
    dist = np.sqrt(_dot(dp, dp))
    ok = dist != 0.0
    ds = np.where(ok, dist, 1.0)
    a1 = _dot(n1, dp) / ds
    a2 = _dot(n2, dp) / ds
    with np.errstate(invalid="ignore"):
        swap = np.arccos(np.abs(a1)) > np.arccos(np.abs(a2))
    n1c = np.where(swap[:, None], n2, n1)
    n2c = np.where(swap[:, None], n1, n2)
    dp = np.where(swap[:, None], -dp, dp)
    f2 = np.where(swap, -a2, a1)
    v = _cross(dp, n1c)
    vn = np.sqrt(_dot(v, v))
    ok &= vn != 0.0
    v = v / np.where(vn != 0, vn, 1.0)[:, None]
    w = _cross(n1c, v)
    f1 = _dot(v, n2c)
    f0 = np.arctan2(_dot(w, n2c), _dot(n1c, n2c))
    z = np.zeros_like(f0)
    return np.where(ok, f0, z), np.where(ok, f1, z), np.where(ok, f2, z)


def _bin(x):
    return np.clip(np.floor(x).astype(np.int64), 0, 10)


def compute_fpfh(points, normals, idx, d2, cnt):
    """FPFH-33 [n,33] float64 from hybrid-search results of the same cloud (o3d.registration.compute_fpfh_feature,
    demo_registration.py:41)."""
    points = np.asarray(points, np.float64)
    n, max_nn = idx.shape
    rows = np.arange(n)
    spfh = np.zeros((n, 33))
    incr = 100.0 / np.maximum(cnt - 1, 1).astype(np.float64)
    for k in range(1, max_nn):
        valid = k < cnt
        j = np.where(valid, idx[:, k], 0)
        f0, f1, f2 = _pair_features(points, normals, points[j], normals[j])
        add = np.where(valid, incr, 0.0)
        spfh[rows, _bin(11 * (f0 + np.pi) / (2.0 * np.pi))] += add
        spfh[rows, 11 + _bin(11 * (f1 + 1.0) * 0.5)] += add
        spfh[rows, 22 + _bin(11 * (f2 + 1.0) * 0.5)] += add
    fp = np.zeros((n, 33))
    for k in range(1, max_nn):
        valid = (k < cnt) & (d2[:, k] != 0.0)
        j = np.where(valid, idx[:, k], 0)
        dist = np.where(valid, d2[:, k], 1.0)
        fp += np.where(valid[:, None], spfh[j] / dist[:, None], 0.0)
    out = np.empty((n, 33))
    for g in range(3):
        blk = fp[:, 11 * g:11 * g + 11]
        s = np.zeros(n)
        for b in range(11):
            s = s + blk[:, b]
        scale = np.where(s != 0.0, 100.0 / np.where(s != 0.0, s, 1.0), 0.0)
        out[:, 11 * g:11 * g + 11] = blk * scale[:, None] + spfh[:, 11 * g:11 * g + 11]
    out[cnt <= 1] = 0.0
    return out


# ----------------------------------------------------------------------------------------------- the demo path
def extract_fpfh_features(points, downsample):
    """demo_registration.py:37-44 from already-read points: (down-sampled points [m,3], unit-ish features [m,33])."""
    idx, _, cnt = hybrid_search(points, downsample * 2, 30)
    normals = estimate_normals(points, idx, cnt)
    pts, nrm = voxel_down_sample(points, normals, downsample)
    idx, d2, cnt = hybrid_search(pts, downsample * 5, 100)
    feat = compute_fpfh(pts, nrm, idx, d2, cnt)
    norm = np.sqrt((feat * feat).sum(1, keepdims=True))
    return pts, feat / (norm + 1e-6)


def ordered_dot(a, b):
    """a @ b.T with every dot product summed in ascending dimension order (no BLAS reordering)."""
    acc = np.zeros((a.shape[0], b.shape[0]))
    for d in range(a.shape[1]):
        acc += a[:, d, None] * b[None, :, d]
    return acc


def match_and_centre(src_pts, src_feat, tgt_pts, tgt_feat):
    """demo_registration.py:101-108: source i is paired with argmin_j sqrt(2 - 2 s_i.t_j + 1e-6) (first minimum)."""
    distance = np.sqrt(2 - 2 * ordered_dot(src_feat, tgt_feat) + 1e-6)
    source_idx = np.argmin(distance, axis=1)
    corr = np.concatenate([np.arange(source_idx.shape[0])[:, None], source_idx[:, None]], axis=-1)
    src_keypts = src_pts[corr[:, 0]]
    tgt_keypts = tgt_pts[corr[:, 1]]
    corr_pos = np.concatenate([src_keypts, tgt_keypts], axis=-1)
    corr_pos = corr_pos - corr_pos.mean(0)
    return corr, src_keypts, tgt_keypts, corr_pos
