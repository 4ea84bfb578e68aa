"""Shared helpers of the GPU parity tests: tolerances (BASELINE.json north_star) and
tie-aware comparators for integer outputs that hinge on floating-point thresholds."""
import torch

from oracle import restate, synth

ROT_TOL_DEG = 0.01      # rotation <= 0.01 deg
TRANS_TOL = 1e-4        # translation <= 1e-4 m
CONF_TOL = 1e-4         # confidences <= 1e-4 abs


def golden_model(meta, device="cuda", attn_mode=None, mlp_mode=None):
    from pointdsc_b200 import PointDSC
    fam = synth.CONFIGS[meta["family"]]
    sd = synth.make_state_dict(meta["weight_seed"], num_layers=meta["num_layers"], sigma_d=fam["sigma_d"])
    m = PointDSC(**meta["ctor"]).eval()
    m.load_state_dict(sd, strict=True)
    m = m.to(device)
    if attn_mode is not None:
        m.attn_mode = attn_mode
    if mlp_mode is not None:
        m.mlp_mode = mlp_mode
    return m, sd


def point_distances(T, src, tgt):
    """fp64 |T src - tgt| for T [4,4], src/tgt [N,3] (CPU)."""
    T, src, tgt = T.double().cpu(), src.double().cpu(), tgt.double().cpu()
    return torch.linalg.norm(src @ T[:3, :3].T + T[:3, 3] - tgt, dim=-1)


def assert_labels_match(labels_a, labels_b, T, src, tgt, thr, band=2e-6, what="labels"):
    """0/1 masks must agree except where the distance sits within `band` of the threshold
    (an exact-threshold tie decided by fp32 summation order); ties are reported."""
    a, b = labels_a.cpu().flatten(), labels_b.cpu().flatten()
    diff = torch.nonzero(a != b).flatten()
    if diff.numel() == 0:
        return 0
    d = point_distances(T, src, tgt)[diff]
    scale = max(1.0, float(src.abs().max()))
    bad = (d - thr).abs() > band * scale
    assert not bad.any(), f"{what}: {int(bad.sum())} mismatches away from the threshold (d={d[bad][:5]})"
    print(f"[tie] {what}: {diff.numel()} exact-threshold ties at indices {diff[:8].tolist()}")
    return diff.numel()


def assert_counts_match(counts_a, counts_b, seed_trans, src, tgt, thr, band=2e-6, what="inlier_count"):
    """Per-seed integer inlier counts: equal, or the difference is explained by points within
    `band` of the threshold under that seed's transform."""
    a, b = counts_a.cpu().flatten().long(), counts_b.cpu().flatten().long()
    diff = torch.nonzero(a != b).flatten()
    scale = max(1.0, float(src.abs().max()))
    for s in diff.tolist():
        d = point_distances(seed_trans.reshape(-1, 4, 4)[s], src, tgt)
        near = int(((d - thr).abs() <= band * scale).sum())
        assert abs(int(a[s] - b[s])) <= near, f"{what}[{s}]: {int(a[s])} vs {int(b[s])}, only {near} near-threshold points"
    if diff.numel():
        print(f"[tie] {what}: {diff.numel()} seeds differ by exact-threshold ties")
    return diff.numel()


def trans_errors(Ta, Tb):
    Ta, Tb = Ta.cpu(), Tb.cpu()
    re = restate.rotation_error_deg(Ta[..., :3, :3], Tb[..., :3, :3])
    te = restate.translation_error(Ta[..., :3, 3], Tb[..., :3, 3])
    return re, te


def assert_trans_close(Ta, Tb, what="final_trans", rot=ROT_TOL_DEG, trans=TRANS_TOL):
    re, te = trans_errors(Ta, Tb)
    assert float(re.max()) <= rot, f"{what}: rotation error {float(re.max()):.3e} deg > {rot}"
    assert float(te.max()) <= trans, f"{what}: translation error {float(te.max()):.3e} m > {trans}"
    return float(re.max()), float(te.max())
