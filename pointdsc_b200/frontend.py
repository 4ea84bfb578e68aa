"""Demo front end on the GPU: PLY -> normals -> voxel down-sample -> FPFH-33 -> putative
correspondences, i.e. everything `demo_registration.py` does before `model(data)` with `--descriptor fpfh`, without
open3d.

Host-side mirror of the reference helpers, same names and argument meaning:

  estimate_normal(points, radius, max_nn)         utils/pointcloud.py:20-21
  extract_fpfh_features(pcd_path, downsample, device)   demo_registration.py:37-44
  build_correspondences(...)                       demo_registration.py:101-116 (matching, centring, the data dict)
  register(model, pcd1, pcd2, downsample)          the whole demo minus the two visualiser calls (no graphics here)

Everything is float64 on the device like open3d / numpy in the reference; there is no CPU path (a CUDA device and the
built library are required).  The only host work is parsing the PLY header and handing the vertex bytes to the device.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _cabi

_PLY_TYPES = {"char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2", "ushort": "u2",
              "uint16": "u2", "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4", "float": "f4", "float32": "f4",
              "double": "f8", "float64": "f8"}


def read_point_cloud(path, device="cuda"):
    """Vertex positions of a PLY file as a float64 [n,3] CUDA tensor (o3d.io.read_point_cloud(...).points,
    demo_registration.py:38).  binary_little_endian / binary_big_endian / ascii; other vertex properties are skipped."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError(f"{path}: not a PLY file")
        fmt, n, props, in_vertex = None, None, [], False
        while True:
            line = f.readline()
            if not line:
                raise ValueError(f"{path}: PLY header not terminated")
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
                    raise ValueError(f"{path}: list properties on vertices are not supported")
                props.append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if fmt is None or n is None or not all(c in [p[0] for p in props] for c in "xyz"):
            raise ValueError(f"{path}: PLY header has no format / vertex x,y,z")
        names = [p[0] for p in props]
        if fmt == "ascii":
            rows = np.loadtxt(f, max_rows=n, ndmin=2)
            xyz = np.stack([rows[:, names.index(c)] for c in "xyz"], 1).astype(np.float64)
        else:
            end = "<" if fmt == "binary_little_endian" else ">"
            dt = np.dtype([(nm, end + ty) for nm, ty in props])
            rec = np.frombuffer(f.read(n * dt.itemsize), dtype=dt, count=n)
            xyz = np.stack([rec[c] for c in "xyz"], 1)
    t = torch.from_numpy(np.ascontiguousarray(xyz)).to(device)
    return t.double().contiguous()


def _f64(x, device=None):
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(x)
    if device is not None:
        x = x.to(device)
    return x.double().contiguous()


def estimate_normal(points, radius=0.06, max_nn=30):
    """Unoriented normals fp64 [n,3] from the max_nn nearest points within radius (utils/pointcloud.py:20-21)."""
    return _cabi.hybrid_search(_f64(points), radius, max_nn, want_lists=False, want_normals=True)[3]


def voxel_down_sample(points, normals, voxel_size):
    """(points [m,3], normals [m,3] or None): PointCloud.voxel_down_sample (demo_registration.py:40)."""
    return _cabi.voxel_down_sample(_f64(points), None if normals is None else _f64(normals), voxel_size)


def compute_fpfh_feature(points, normals, radius, max_nn=100):
    """FPFH-33 fp64 [n,33] (row per point; open3d's Feature.data is the transpose), demo_registration.py:41."""
    points = _f64(points)
    idx, d2, cnt, _ = _cabi.hybrid_search(points, radius, max_nn)
    return _cabi.fpfh(points, _f64(normals), idx, d2, cnt)


def extract_fpfh_features(pcd_path, downsample, device="cuda"):
    """demo_registration.py:37-44: (raw points [n,3], down-sampled points [m,3], L2-normalised FPFH [m,33]), fp64 CUDA.
    `pcd_path` may also be an [n,3] array / tensor of already-read points."""
    raw = read_point_cloud(pcd_path, device) if isinstance(pcd_path, (str, bytes)) else _f64(pcd_path, device)
    normals = estimate_normal(raw, radius=downsample * 2)
    pts, nrm = voxel_down_sample(raw, normals, downsample)
    feat = compute_fpfh_feature(pts, nrm, radius=downsample * 5, max_nn=100)
    feat = feat / (feat.norm(dim=1, keepdim=True) + 1e-6)
    return raw, pts, feat


def build_correspondences(src_pts, src_features, tgt_pts, tgt_features):
    """demo_registration.py:101-116: nearest target descriptor of every source point, key points, centred corr_pos and
    the `data` dict the model takes (fp32, batch of one)."""
    source_idx = _cabi.nn_match_f64(_f64(src_features), _f64(tgt_features))
    corr = torch.stack([torch.arange(source_idx.shape[0], device=source_idx.device), source_idx], dim=-1)
    src_keypts = src_pts[corr[:, 0]]
    tgt_keypts = tgt_pts[corr[:, 1]]
    corr_pos = torch.cat([src_keypts, tgt_keypts], dim=-1)
    corr_pos = corr_pos - corr_pos.mean(0)
    data = {"corr_pos": corr_pos[None].float(), "src_keypts": src_keypts[None].float(),
            "tgt_keypts": tgt_keypts[None].float(), "testing": True}
    return corr, data


def register(model, pcd1, pcd2, downsample=0.05, device="cuda"):
    """The demo end to end (demo_registration.py:93-118 without the two draw calls): returns the model's result dict
    plus the correspondences and down-sampled clouds."""
    _, src_pts, src_feat = extract_fpfh_features(pcd1, downsample, device)
    _, tgt_pts, tgt_feat = extract_fpfh_features(pcd2, downsample, device)
    corr, data = build_correspondences(src_pts, src_feat, tgt_pts, tgt_feat)
    res = dict(model(data))
    res.update(corr=corr, src_pts=src_pts, tgt_pts=tgt_pts)
    return res
