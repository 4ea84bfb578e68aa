"""Stand-in for compute-sanitizer (closed on this GPU pool, profiles/r02_sanitizer_closed.txt): the whole forward and the
attention entry points run on a workspace that is (a) poisoned with 0xFF bytes (NaN as fp32, -1 as integers) before every
call and (b) surrounded by guard bands.  A kernel that reads scratch it did not write produces different results (the
comparison with a plain run is bit for bit); one that writes outside its workspace breaks a guard band."""
import pytest
import torch

from oracle import synth

pytestmark = pytest.mark.gpu
GUARD = 1 << 16


class GuardedWorkspace:
    def __init__(self):
        self.buf, self.n = None, 0
        self.calls = 0

    def get(self, nbytes, device):
        nbytes = (max(int(nbytes), 256) + 1023) // 1024 * 1024
        if self.buf is None or self.buf.numel() < nbytes + 2 * GUARD or self.buf.device != device:
            self.buf = torch.empty(nbytes + 2 * GUARD, dtype=torch.uint8, device=device)
        else:
            self.check()
        self.buf.fill_(0xFF)
        self.n = nbytes
        self.calls += 1
        return self.buf[GUARD:GUARD + nbytes]

    def check(self):
        assert bool((self.buf[:GUARD] == 0xFF).all()), "guard band below the workspace was written"
        assert bool((self.buf[GUARD + self.n:] == 0xFF).all()), "guard band above the workspace was written"


@pytest.mark.parametrize("fam,N,B,L", [("3dmatch", 700, 3, 3), ("kitti", 1300, 2, 2), ("3dmatch", 257, 1, 2)])
def test_forward_on_poisoned_guarded_workspace_is_bit_identical(fam, N, B, L):
    from pointdsc_b200 import PointDSC
    cfg = synth.CONFIGS[fam]
    sd = synth.make_state_dict(99, num_layers=L, sigma_d=cfg["sigma_d"])
    model = PointDSC(num_layers=L, inlier_threshold=cfg["inlier_threshold"], sigma_d=cfg["sigma_d"], nms_radius=cfg["nms_radius"]).eval()
    model.load_state_dict(sd, strict=True)
    model = model.cuda()
    batch = synth.make_batch(list(range(40, 40 + B)), N, **cfg)
    data = {k: batch[k].cuda() for k in ("corr_pos", "src_keypts", "tgt_keypts")}
    data["testing"] = True
    plain = model(data)
    torch.cuda.synchronize()
    model._ws = GuardedWorkspace()
    guarded = model(data)
    torch.cuda.synchronize()
    model._ws.check()
    assert model._ws.calls >= L + 3          # qkv_pack per layer + pick_seeds + seed_knn + seed_hypotheses
    for k in ("final_trans", "final_labels", "confidence", "seeds", "knn_idx", "seed_trans", "inlier_count", "corr_features"):
        assert torch.equal(plain[k], guarded[k]), k


@pytest.mark.parametrize("N", [100, 700, 1300])
def test_attention_on_poisoned_guarded_workspace(N):
    """pdsc_nonlocal_attn (prep + main kernel, stream-K partials at these sizes) on a poisoned workspace."""
    from pointdsc_b200 import _cabi
    g = torch.Generator().manual_seed(N)
    C, B = 128, 2
    qkv = torch.randn(B, N, 3 * C, generator=g).cuda()
    src = torch.rand(B, N, 3, generator=g).cuda()
    tgt = (src + 0.05 * torch.randn(B, N, 3, generator=g).cuda())
    plain = _cabi.nonlocal_attn(qkv[:, :, :C], qkv[:, :, C:2 * C], qkv[:, :, 2 * C:], src, tgt, 0.1, mode=_cabi.ATTN_SPLIT16)
    ws = GuardedWorkspace()
    out = _cabi.nonlocal_attn(qkv[:, :, :C], qkv[:, :, C:2 * C], qkv[:, :, 2 * C:], src, tgt, 0.1, mode=_cabi.ATTN_SPLIT16, ws=ws)
    torch.cuda.synchronize()
    ws.check()
    assert torch.isfinite(out).all()
    assert torch.equal(plain, out)
