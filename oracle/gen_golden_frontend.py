"""Writes tests/golden/front_demo_crop.npz: a crop of the reference's two demo clouds (demo_data/cloud_bin_{0,1}.ply,
BASELINE config 1) and what oracle/restate_frontend.py makes of it.  Run in the build container (reads /root/reference):

    python -m oracle.gen_golden_frontend

open3d is not installed, so these vectors pin the restatement against itself (regression) and give the GPU tests a real
scan to work on; they are NOT outputs of the reference's open3d calls (see the header of restate_frontend.py).
"""
import os

import numpy as np

from . import restate_frontend as F

REF = os.environ.get("POINTDSC_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "front_demo_crop.npz")
DOWNSAMPLE = 0.05   # snapshot/PointDSC_3DMatch_release/config.json "downsample"


def crop(points, n_keep):
    """The n_keep points nearest to the cloud's median point (a compact, dense piece of the scan), input order kept."""
    c = np.median(points, 0)
    d = ((points - c) ** 2).sum(1)
    keep = np.sort(np.argsort(d, kind="stable")[:n_keep])
    return points[keep]


def main():
    out = {}
    for name, f in (("src", "cloud_bin_0.ply"), ("tgt", "cloud_bin_1.ply")):
        pts = crop(F.read_ply_xyz(os.path.join(REF, "demo_data", f)), 24000)
        out[f"{name}_raw"] = pts.astype(np.float32)          # the PLY stores float32: lossless
        idx, d2, cnt = F.hybrid_search(pts, DOWNSAMPLE * 2, 30)
        nrm = F.estimate_normals(pts, idx, cnt)
        dpts, dnrm = F.voxel_down_sample(pts, nrm, DOWNSAMPLE)
        idx2, d22, cnt2 = F.hybrid_search(dpts, DOWNSAMPLE * 5, 100)
        feat = F.compute_fpfh(dpts, dnrm, idx2, d22, cnt2)
        sel = np.arange(0, len(pts), 37)
        out[f"{name}_sel"] = sel.astype(np.int32)
        out[f"{name}_sel_idx"] = idx[sel]
        out[f"{name}_sel_normals"] = nrm[sel]
        out[f"{name}_down_pts"] = dpts
        out[f"{name}_down_normals"] = dnrm
        out[f"{name}_down_cnt"] = cnt2
        out[f"{name}_fpfh"] = feat.astype(np.float32)
        print(name, pts.shape, "->", dpts.shape, "mean neighbours", cnt.mean(), cnt2.mean())
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT) // 1024, "KiB")


if __name__ == "__main__":
    main()
