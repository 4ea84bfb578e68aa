"""GPU parity of the ICP hand-off (pointdsc_b200/benchmark_utils.py, csrc/icp.cu) against the CPU
restatement oracle/restate_icp.py: associations and fitness exact, transforms within the forward's tolerances."""
import numpy as np
import pytest
import torch

from oracle import restate_icp as I, synth
from parity import assert_trans_close

pytestmark = pytest.mark.gpu


def _perturb(T, seed, angle_deg=1.0, shift=0.02):
    rng = np.random.default_rng(seed)
    ax = rng.normal(size=3); ax /= np.linalg.norm(ax)
    a = np.deg2rad(angle_deg)
    K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    D = np.eye(4)
    D[:3, :3] = np.eye(3) + np.sin(a) * K + (1 - np.cos(a)) * K @ K
    D[:3, 3] = rng.normal(size=3) * shift
    return torch.from_numpy(D @ T.double().numpy()).float()


def _check(src, tgt, T0, max_dist, **kw):
    from pointdsc_b200 import _cabi
    out, stats, iters, corr = _cabi.icp_refine(src.cuda(), tgt.cuda(), T0.cuda(), max_dist=max_dist, want_corr=True, **kw)
    for b in range(src.shape[0]):
        Tr, fit, rmse, it, idx = I.icp_point_to_point(src[b].numpy(), tgt[b].numpy(), T0[b].numpy(), max_dist=max_dist, **kw)
        same = (corr[b].cpu().numpy() == idx).mean()
        print(f"pair {b}: updates {int(iters[b])} vs {it}, fitness {float(stats[b, 0]):.4f} vs {fit:.4f}, "
              f"rmse {float(stats[b, 1]):.6f} vs {rmse:.6f}, associations identical {100 * same:.2f}%")
        assert int(iters[b]) == it
        assert same == 1.0
        assert float(stats[b, 0]) == fit                       # integer count / N
        assert abs(float(stats[b, 1]) - rmse) <= 1e-9
        assert_trans_close(out[b:b + 1], torch.from_numpy(Tr).float()[None], what=f"icp pair {b}")
    return out, stats, iters


@pytest.mark.parametrize("fam,N,ratio", [("3dmatch", 1000, 0.6), ("3dmatch", 5000, 0.1), ("kitti", 3000, 0.3)])
def test_icp_refine_vs_oracle(fam, N, ratio):
    cfg = dict(synth.CONFIGS[fam], inlier_ratio=ratio)
    batch = synth.make_batch([31, 32, 33], N, **cfg)
    T0 = torch.stack([_perturb(batch["gt_trans"][b], b, angle_deg=0.5 * (b + 1)) for b in range(3)])
    _check(batch["src_keypts"], batch["tgt_keypts"], T0, max_dist=cfg["inlier_threshold"])


def test_icp_refine_improves_a_perturbed_ground_truth():
    cfg = dict(synth.CONFIGS["3dmatch"], inlier_ratio=0.8, noise=0.005)
    p = synth.make_pair(77, 4000, **cfg)
    T0 = _perturb(p["gt_trans"], 5, angle_deg=2.0, shift=0.03)[None]
    out, stats, iters = _check(p["src_keypts"], p["tgt_keypts"], T0, max_dist=0.1)
    from oracle import restate
    before = float(restate.rotation_error_deg(T0[0, :3, :3], p["gt_trans"][:3, :3]))
    after = float(restate.rotation_error_deg(out[0, :3, :3].cpu(), p["gt_trans"][:3, :3]))
    print(f"rotation error {before:.3f} -> {after:.4f} deg in {int(iters[0])} updates")
    assert after < 0.1 * before


def test_icp_refine_edge_cases():
    from pointdsc_b200 import benchmark_utils
    rng = np.random.default_rng(3)
    src = torch.from_numpy(rng.uniform(0, 1, (2, 300, 3))).float()
    tgt = src.clone()
    tgt[1] += 50.0                                               # pair 1: no target within reach
    T0 = torch.eye(4)[None].repeat(2, 1, 1)
    _check(src, tgt, T0, max_dist=0.1)
    _check(src, tgt, T0, max_dist=0.1, max_iter=0)               # no update at all
    one = torch.zeros(1, 1, 3)
    _check(one, one + 0.01, torch.eye(4)[None], max_dist=0.1)     # a single point
    out, info = benchmark_utils.icp_refine(src.cuda(), tgt.cuda(), T0.cuda(), return_stats=True)
    assert out.shape == (2, 4, 4) and out.is_cuda and float(info["fitness"][0]) == 1.0 and float(info["fitness"][1]) == 0.0
    assert torch.equal(out[1].cpu(), torch.eye(4))
