"""Times the demo front end on the GPU, stage by stage, on a synthetic indoor scan of the demo clouds'
size (258 342 points, ~5 mm spacing), and the CPU restatement on a bounded sample of the same cloud.
python scripts/bench_frontend.py > profiles/r01_frontend.json"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from pointdsc_b200 import _cabi, frontend
from oracle import restate_frontend as F


def synthetic_scan(n, seed):
    """Floor, two walls and a few boxes in a 2.7 x 1.6 x 3 m volume (the demo clouds' extent), 1 mm noise."""
    rng = np.random.default_rng(seed)
    parts = []
    faces = [((0, 0, 0), (2.7, 0, 0), (0, 1.6, 0)), ((0, 0, 0), (2.7, 0, 0), (0, 0, 3.0)), ((0, 0, 0), (0, 1.6, 0), (0, 0, 3.0))]
    for b in range(4):
        o = rng.uniform([0.3, 0.0, 0.3], [2.0, 0.0, 2.2]); s = rng.uniform(0.3, 0.7, 3)
        faces += [((o[0], s[1], o[2]), (s[0], 0, 0), (0, 0, s[2])), (tuple(o), (s[0], 0, 0), (0, s[1], 0)), (tuple(o), (0, s[1], 0), (0, 0, s[2]))]
    area = np.array([np.linalg.norm(np.cross(u, v)) for _, u, v in faces])
    counts = np.floor(n * area / area.sum()).astype(int); counts[0] += n - counts.sum()
    for (o, u, v), k in zip(faces, counts):
        ab = rng.uniform(0, 1, (k, 2))
        nrm = np.cross(u, v); nrm = nrm / np.linalg.norm(nrm)
        parts.append(np.asarray(o) + ab[:, :1] * np.asarray(u) + ab[:, 1:] * np.asarray(v) + rng.normal(0, 1e-3, (k, 1)) * nrm)
    return np.vstack(parts).astype(np.float32).astype(np.float64)[rng.permutation(n)]


def gpu_ms(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


n = 258342
src = torch.from_numpy(synthetic_scan(n, 0)).cuda()
tgt = torch.from_numpy(synthetic_scan(n, 1)).cuda()
out = {"workload": f"two synthetic indoor scans of {n} points (size of demo_data/cloud_bin_0.ply), downsample 0.05, "
                   "normals radius 0.1 / 30 nn, FPFH radius 0.25 / 100 nn, float64"}
ms_n = gpu_ms(lambda: frontend.estimate_normal(src, radius=0.1, max_nn=30))
nrm = frontend.estimate_normal(src, radius=0.1, max_nn=30)
ms_v = gpu_ms(lambda: frontend.voxel_down_sample(src, nrm, 0.05))
dp, dn = frontend.voxel_down_sample(src, nrm, 0.05)
ms_s = gpu_ms(lambda: _cabi.hybrid_search(dp, 0.25, 100))
idx, d2, cnt, _ = _cabi.hybrid_search(dp, 0.25, 100)
ms_f = gpu_ms(lambda: _cabi.fpfh(dp, dn, idx, d2, cnt))
_, sp, sf = frontend.extract_fpfh_features(src, 0.05)
_, tp, tf = frontend.extract_fpfh_features(tgt, 0.05)
ms_m = gpu_ms(lambda: frontend.build_correspondences(sp, sf, tp, tf))
ms_all = gpu_ms(lambda: (frontend.extract_fpfh_features(src, 0.05), frontend.extract_fpfh_features(tgt, 0.05),
                         frontend.build_correspondences(sp, sf, tp, tf)), reps=5)
# distance tests of the raw-cloud search: candidates scanned per query (27-cell neighbourhood), counted on the host grid
h = 0.1
pts = src.cpu().numpy()
cell = np.floor((pts - pts.min(0)) / h).astype(np.int64)
dims = cell.max(0) + 1
key = (cell[:, 2] * dims[1] + cell[:, 1]) * dims[0] + cell[:, 0]
occ = np.zeros(dims[::-1], np.int64)
np.add.at(occ.reshape(-1), key, 1)
pad = np.pad(occ, 1)
neigh = sum(pad[1 + a:pad.shape[0] - 1 + a, 1 + b:pad.shape[1] - 1 + b, 1 + c:pad.shape[2] - 1 + c]
            for a in (-1, 0, 1) for b in (-1, 0, 1) for c in (-1, 0, 1))
tests = float((occ * neigh).sum())
out["gpu_ms"] = {"estimate_normal (search + covariance + eigenvector, 258k points)": ms_n, "voxel_down_sample": ms_v,
                 f"hybrid_search 0.25/100 on {dp.shape[0]} down-sampled points": ms_s, "fpfh (spfh + weighting)": ms_f,
                 f"matching + centring ({sp.shape[0]} x {tp.shape[0]} x 33)": ms_m, "both clouds end to end": ms_all}
out["pairs_per_s_front_end"] = 1e3 / ms_all
out["normals_search"] = {"distance_tests": tests, "tests_per_query": tests / n, "fp64_ops_per_test": 8,
                         "achieved_fp64_TFLOPs": tests * 8 / ms_n / 1e9,
                         "note": "bound: FP64 ALU (8 double ops per candidate, 64 lanes/clk/SM on B200); algorithmic HBM bytes "
                                 f"{n * 24 * 2} (points in, normals out) are negligible"}
# CPU restatement on a bounded sample: 20 000 queries of the same cloud for the normals, everything after on the full down-sampled cloud
t0 = time.perf_counter()
sub = pts[:20000]
ridx, rd2, rcnt = F.hybrid_search(sub, 0.1, 30)
F.estimate_normals(sub, ridx, rcnt)
cpu_norm = (time.perf_counter() - t0) * n / 20000
t0 = time.perf_counter()
dpc, dnc = F.voxel_down_sample(pts, nrm.cpu().numpy(), 0.05)
ridx, rd2, rcnt = F.hybrid_search(dpc, 0.25, 100)
F.compute_fpfh(dpc, dnc, ridx, rd2, rcnt)
cpu_rest = time.perf_counter() - t0
out["cpu_port"] = {"normals_s_per_cloud_extrapolated_from_20000_queries": cpu_norm, "voxel+fpfh_s_per_cloud": cpu_rest,
                   "s_per_pair": 2 * (cpu_norm + cpu_rest), "threads": os.cpu_count(),
                   "kind": "port (oracle/restate_frontend.py: scipy cKDTree + numpy)"}
out["speedup_vs_cpu_port"] = out["cpu_port"]["s_per_pair"] / (ms_all / 1e3)
print(json.dumps(out, indent=1))
