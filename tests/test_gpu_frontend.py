"""GPU parity of the demo front end (pointdsc_b200/frontend.py, csrc/frontend.cu) against the CPU
restatement oracle/restate_frontend.py: neighbour lists and voxel assignments bit-exact, float64 outputs to 1e-9, on
synthetic clouds and on the committed crop of the reference's demo clouds; then the demo end to end (BASELINE
config 1 shape: two clouds -> FPFH -> correspondences -> PointDSC.forward)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN_DIR
from oracle import restate, restate_frontend as F, synth
from parity import CONF_TOL, assert_labels_match, assert_trans_close

pytestmark = pytest.mark.gpu

FP_TOL = 1e-9      # float64 outputs (normals, voxel means, FPFH bins, matching distances): absolute tolerance


def _golden():
    return np.load(os.path.join(GOLDEN_DIR, "front_demo_crop.npz"))


def _cloud(kind, n, seed):
    rng = np.random.default_rng(seed)
    if kind == "uniform":
        return rng.uniform(0, 1, (n, 3))
    if kind == "float32_lattice":        # coordinates on a coarse lattice: many exact distance ties and duplicates
        return (rng.integers(0, 12, (n, 3)) * 0.05).astype(np.float32).astype(np.float64)
    if kind == "room":                    # three noisy walls, like a scan
        k = n // 3
        a = np.c_[rng.uniform(0, 2, k), rng.uniform(0, 2, k), rng.normal(0, 2e-3, k)]
        b = np.c_[rng.uniform(0, 2, k), rng.normal(0, 2e-3, k), rng.uniform(0, 2, k)]
        c = np.c_[rng.normal(0, 2e-3, n - 2 * k), rng.uniform(0, 2, n - 2 * k), rng.uniform(0, 2, n - 2 * k)]
        return np.vstack([a, b, c]).astype(np.float32).astype(np.float64)
    raise ValueError(kind)


@pytest.mark.parametrize("kind,n,radius,max_nn", [("uniform", 1, 0.1, 30), ("uniform", 2, 5.0, 30), ("uniform", 700, 0.15, 30),
                                                  ("uniform", 3000, 0.3, 100), ("float32_lattice", 1500, 0.11, 30),
                                                  ("room", 6000, 0.1, 30), ("room", 2000, 0.25, 128),
                                                  # > 1024 candidates inside the radius: the warp-per-query kernel hands these
                                                  # queries to the thread-per-query kernel
                                                  ("uniform", 5000, 0.5, 64), ("float32_lattice", 4000, 0.3, 100),
                                                  ("room", 70000, 0.05, 30)])
def test_hybrid_search_bit_exact(kind, n, radius, max_nn):
    from pointdsc_b200 import _cabi
    pts = _cloud(kind, n, n)
    idx, d2, cnt, _ = _cabi.hybrid_search(torch.from_numpy(pts).cuda(), radius, max_nn)
    ridx, rd2, rcnt = F.hybrid_search(pts, radius, max_nn)
    assert np.array_equal(cnt.cpu().numpy(), rcnt)
    assert np.array_equal(idx.cpu().numpy(), ridx)
    assert np.array_equal(d2.cpu().numpy(), rd2)          # same products, same sums, no FMA: identical bits


def test_hybrid_search_huge_extent_grows_the_cells():
    """Two far-apart clusters: the search grid would need more cells than the workspace holds at cell = radius."""
    from pointdsc_b200 import _cabi
    rng = np.random.default_rng(5)
    pts = np.vstack([rng.uniform(0, 1, (300, 3)), rng.uniform(0, 1, (300, 3)) + 4000.0])
    idx, d2, cnt, _ = _cabi.hybrid_search(torch.from_numpy(pts).cuda(), 0.2, 20)
    ridx, rd2, rcnt = F.hybrid_search(pts, 0.2, 20)
    assert np.array_equal(idx.cpu().numpy(), ridx) and np.array_equal(cnt.cpu().numpy(), rcnt)


@pytest.mark.parametrize("name", ["src", "tgt"])
def test_normals_on_demo_crop(name):
    from pointdsc_b200 import frontend
    g = _golden()
    pts = g[f"{name}_raw"].astype(np.float64)
    nrm = frontend.estimate_normal(torch.from_numpy(pts).cuda(), radius=0.1, max_nn=30).cpu().numpy()
    sel = g[f"{name}_sel"]
    err = np.abs(nrm[sel] - g[f"{name}_sel_normals"]).max()
    print(f"{name}: max |normal - oracle| = {err:.2e} over {len(sel)} stored points")
    assert err <= FP_TOL
    # and the whole crop against the live oracle
    ridx, rd2, rcnt = F.hybrid_search(pts, 0.1, 30)
    ref = F.estimate_normals(pts, ridx, rcnt)
    assert np.abs(nrm - ref).max() <= FP_TOL
    assert np.allclose(np.linalg.norm(nrm, axis=1), 1.0, atol=1e-12)


def test_normals_degenerate_inputs():
    from pointdsc_b200 import frontend
    line = np.c_[np.linspace(0, 1, 50), np.zeros(50), np.zeros(50)]           # rank-1 covariance
    same = np.tile([[0.3, 0.2, 0.1]], (40, 1))                                  # zero covariance -> (0,0,1)
    lonely = np.array([[0.0, 0, 0], [10.0, 0, 0], [20.0, 0, 0]])               # < 3 neighbours -> (0,0,1)
    for pts in (line, same, lonely):
        got = frontend.estimate_normal(torch.from_numpy(pts).cuda(), radius=0.5, max_nn=30).cpu().numpy()
        ridx, _, rcnt = F.hybrid_search(pts, 0.5, 30)
        ref = F.estimate_normals(pts, ridx, rcnt)
        assert np.abs(got - ref).max() <= FP_TOL
    assert (got == [0, 0, 1]).all()


@pytest.mark.parametrize("kind,n,voxel", [("uniform", 1, 0.05), ("uniform", 5000, 0.1), ("room", 9000, 0.05),
                                          ("float32_lattice", 4000, 0.05)])
def test_voxel_down_sample_vs_oracle(kind, n, voxel):
    from pointdsc_b200 import frontend
    pts = _cloud(kind, n, 100 + n)
    nrm = np.random.default_rng(n).normal(size=(n, 3))
    gp, gn = frontend.voxel_down_sample(torch.from_numpy(pts).cuda(), torch.from_numpy(nrm).cuda(), voxel)
    rp, rn = F.voxel_down_sample(pts, nrm, voxel)
    assert gp.shape == rp.shape                                     # same occupied voxels ...
    assert np.array_equal(gp.cpu().numpy(), rp)                      # ... same order, same sums in the same order
    assert np.abs(gn.cpu().numpy() - rn).max() <= 1e-15
    gp2, gn2 = frontend.voxel_down_sample(torch.from_numpy(pts).cuda(), None, voxel)
    assert gn2 is None and torch.equal(gp2, gp)


def test_voxel_down_sample_grid_larger_than_workspace_retries():
    from pointdsc_b200 import _cabi
    rng = np.random.default_rng(6)
    pts = np.vstack([rng.uniform(0, 1, (200, 3)), rng.uniform(0, 1, (200, 3)) + [20.0, 10.0, 5.0]])
    gp, _ = _cabi.voxel_down_sample(torch.from_numpy(pts).cuda(), None, 0.05, cells=4096)    # needs ~9e6 cells
    rp, _ = F.voxel_down_sample(pts, None, 0.05)
    assert np.array_equal(gp.cpu().numpy(), rp)
    far = np.vstack([pts, [[1e6, 1e6, 1e6]]])
    with pytest.raises(RuntimeError):
        _cabi.voxel_down_sample(torch.from_numpy(far).cuda(), None, 0.05)


@pytest.mark.parametrize("name", ["src", "tgt"])
def test_fpfh_on_demo_crop(name):
    from pointdsc_b200 import frontend
    g = _golden()
    pts, nrm = g[f"{name}_down_pts"], g[f"{name}_down_normals"]
    got = frontend.compute_fpfh_feature(torch.from_numpy(pts).cuda(), torch.from_numpy(nrm).cuda(), radius=0.25,
                                        max_nn=100).cpu().numpy()
    ridx, rd2, rcnt = F.hybrid_search(pts, 0.25, 100)
    ref = F.compute_fpfh(pts, nrm, ridx, rd2, rcnt)
    bad = np.abs(got - ref) > FP_TOL
    # a bin decision (floor of an angle expression) can sit within an ulp of a bin edge, where device and host libm may
    # differ: such rows are logged, and must be rare
    rows = np.unique(np.nonzero(bad)[0])
    if len(rows):
        print(f"[tie] fpfh {name}: {len(rows)} of {len(pts)} rows differ by a bin-edge decision")
    assert len(rows) <= max(1, len(pts) // 200)
    assert np.allclose(got, g[f"{name}_fpfh"], atol=2e-4) or len(rows)
    full = rcnt > 1
    for grp in range(3):
        assert np.allclose(got[full, 11 * grp:11 * grp + 11].sum(1), 200.0, atol=1e-9)


def test_fpfh_isolated_points_are_zero_rows():
    from pointdsc_b200 import frontend
    pts = np.array([[0.0, 0, 0], [0.01, 0, 0], [0, 0.01, 0], [5.0, 5, 5]])
    nrm = np.tile([[0, 0, 1.0]], (4, 1))
    got = frontend.compute_fpfh_feature(torch.from_numpy(pts).cuda(), torch.from_numpy(nrm).cuda(), radius=0.1).cpu().numpy()
    ridx, rd2, rcnt = F.hybrid_search(pts, 0.1, 100)
    assert np.abs(got - F.compute_fpfh(pts, nrm, ridx, rd2, rcnt)).max() <= FP_TOL
    assert (got[3] == 0).all() and got[0].sum() > 0


@pytest.mark.parametrize("Na,Nb,D", [(1, 1, 33), (70, 45, 33), (900, 1300, 33), (33, 200, 64), (257, 65, 5)])
def test_nn_match_f64_first_minimum(Na, Nb, D):
    from pointdsc_b200 import _cabi
    rng = np.random.default_rng(Na + Nb)
    a = rng.normal(size=(Na, D)); a /= np.linalg.norm(a, axis=1, keepdims=True)
    b = rng.normal(size=(Nb, D)); b /= np.linalg.norm(b, axis=1, keepdims=True)
    if Nb > 10:
        b[9] = b[2]; a[0] = b[2]                                    # equal distances: the first index wins
    idx, val = _cabi.nn_match_f64(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda(), want_values=True)
    dist = np.sqrt(2 - 2 * F.ordered_dot(a, b) + 1e-6)
    assert np.array_equal(idx.cpu().numpy(), np.argmin(dist, axis=1))
    assert np.array_equal(val.cpu().numpy(), dist.min(1))


def test_demo_end_to_end_vs_oracle():
    """BASELINE config 1 shape on the committed crop: clouds -> correspondences -> PointDSC.forward, against the same
    chain on the CPU (restate_frontend + restate.forward_pair).  Seeded weights (the released snapshot is not vendored)."""
    from pointdsc_b200 import PointDSC, frontend
    g = _golden()
    cfg = synth.CONFIGS["3dmatch"]
    L = 3
    sd = synth.make_state_dict(1234, num_layers=L, sigma_d=cfg["sigma_d"])
    model = PointDSC(num_layers=L, inlier_threshold=cfg["inlier_threshold"], sigma_d=cfg["sigma_d"],
                     nms_radius=cfg["nms_radius"]).eval()
    model.load_state_dict(sd, strict=True)
    model = model.cuda()
    src_raw, tgt_raw = g["src_raw"].astype(np.float64), g["tgt_raw"].astype(np.float64)
    res = frontend.register(model, src_raw, tgt_raw, downsample=0.05)

    sp, sf = F.extract_fpfh_features(src_raw, 0.05)
    tp, tf = F.extract_fpfh_features(tgt_raw, 0.05)
    corr, sk, tk, cp = F.match_and_centre(sp, sf, tp, tf)
    assert np.array_equal(res["src_pts"].cpu().numpy(), sp) and np.array_equal(res["tgt_pts"].cpu().numpy(), tp)
    same = (res["corr"].cpu().numpy() == corr).all(1).mean()
    print(f"correspondences identical: {100 * same:.2f}% of {len(corr)}")
    assert same == 1.0
    ref = restate.forward_pair(sd, torch.from_numpy(cp).float()[None], torch.from_numpy(sk).float()[None],
                               torch.from_numpy(tk).float()[None], num_layers=L, inlier_threshold=cfg["inlier_threshold"],
                               nms_radius=cfg["nms_radius"])
    dconf = float((res["confidence"].cpu() - ref["confidence"]).abs().max())
    assert dconf <= CONF_TOL * max(1.0, float(ref["confidence"].abs().max()))
    sk32, tk32 = torch.from_numpy(sk).float()[None], torch.from_numpy(tk).float()[None]
    seeds = res["seeds"].cpu()
    if not torch.equal(seeds, ref["seeds"]):
        # With seeded (untrained) weights on a small crop fewer than S correspondences are positive local maxima; the rest
        # of the seed list has key exactly 0 and the reference's argsort leaves its order open.  The
        # well-defined prefix must agree; the tail is then taken from the GPU run and the remaining stages re-checked.
        n_pos = int((ref["seed_keys"] > 0).sum())
        S = seeds.shape[1]
        assert n_pos < S, "seed lists differ although all keys are positive"
        assert torch.equal(seeds[:, :n_pos], ref["seeds"][:, :n_pos])
        assert set(seeds[0, n_pos:].tolist()).isdisjoint(ref["seeds"][0, :n_pos].tolist())
        print(f"[tie] seeds: {n_pos} of {S} have a positive key; zero-key tail order is unspecified in the reference")
        ref = dict(ref)
        ref.update(restate.seed_hypotheses(seeds, ref["normed_features"], sk32, tk32, sd["sigma"], sd["sigma_spat"], 40, 10,
                                           cfg["inlier_threshold"]))
        ref["final_trans"], _ = restate.post_refinement(ref["initial_trans"], sk32, tk32, cfg["inlier_threshold"])
    assert_trans_close(res["final_trans"], ref["final_trans"])
    assert_labels_match(res["final_labels"][0], ref["final_labels"][0], res["initial_trans"][0], sk32[0], tk32[0],
                        cfg["inlier_threshold"])
