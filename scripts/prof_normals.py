"""One normals pass on a synthetic scan of the demo clouds' size (for ncu captures of hybrid_kernel)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from pointdsc_b200 import frontend, _cabi

rng = np.random.default_rng(0)
n = 258342
k = n // 3
pts = np.vstack([np.c_[rng.uniform(0, 2.7, k), rng.uniform(0, 1.6, k), rng.normal(0, 1e-3, k)],
                 np.c_[rng.uniform(0, 2.7, k), rng.normal(0, 1e-3, k), rng.uniform(0, 3, k)],
                 np.c_[rng.normal(0, 1e-3, n - 2 * k), rng.uniform(0, 1.6, n - 2 * k), rng.uniform(0, 3, n - 2 * k)]])
pts = torch.from_numpy(pts.astype(np.float32).astype(np.float64)).cuda()
for _ in range(3):
    nrm = frontend.estimate_normal(pts, radius=0.1, max_nn=30)
    dp, dn = frontend.voxel_down_sample(pts, nrm, 0.05)
    idx, d2, cnt, _ = _cabi.hybrid_search(dp, 0.25, 100)
    f = _cabi.fpfh(dp, dn, idx, d2, cnt)
torch.cuda.synchronize()
print("ok", dp.shape, float(cnt.float().mean()))
