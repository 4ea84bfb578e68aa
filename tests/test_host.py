"""CPU tests: the C-ABI library loads and exports every symbol the header declares, the
drop-in module has the reference's state_dict layout, and the host-side (torch) part of the
forward -- BN folding, fused q/k/v projection, channel-last MLPs -- reproduces the oracle when
the CUDA attention is replaced by the oracle's (tests only; the product has no such path)."""
import os
import re

import pytest
import torch

from conftest import ROOT, load_golden
from oracle import restate, synth
from pointdsc_b200 import PointDSC, _cabi


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "pointdsc_b200.h")).read()
    declared = set(re.findall(r"\b(pdsc_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    lib = _cabi.load()
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/pointdsc_b200.h but not exported"
    assert declared == set(_cabi.SIGNATURES), declared ^ set(_cabi.SIGNATURES)
    assert lib.pdsc_version() >= 100


def test_product_library_reads_no_tuning_knob():
    """The debug / tuning knobs exist only in libpointdsc_b200_tuning.so (-DPDSC_TUNING); the shipped library does not even
    contain their names, so no environment variable can change its arithmetic or launch shape."""
    blob = open(_cabi.LIB_PATH, "rb").read()
    assert b"PDSC_ATTN_" not in blob and b"PDSC_KNN_" not in blob
    tuning = open(_cabi.TUNING_LIB_PATH, "rb").read()
    for knob in (b"PDSC_ATTN_GRID", b"PDSC_ATTN_CHAIN", b"PDSC_ATTN_CLUSTER", b"PDSC_ATTN_DBG", b"PDSC_ATTN_PV_TERMS",
                 b"PDSC_KNN_FP32"):
        assert knob in tuning, knob
    with _cabi.use_tuning_library() as lib:
        assert _cabi.is_tuning_library() and lib.pdsc_version() >= 100
    assert not _cabi.is_tuning_library()


def test_mode_typos_are_refused(monkeypatch):
    monkeypatch.setenv("POINTDSC_MLP_MODE", "fuse")
    with pytest.raises(ValueError, match="POINTDSC_MLP_MODE"):
        PointDSC(num_layers=1)
    monkeypatch.delenv("POINTDSC_MLP_MODE")
    monkeypatch.setenv("POINTDSC_ATTN_MODE", "split")
    with pytest.raises(ValueError, match="POINTDSC_ATTN_MODE"):
        PointDSC(num_layers=1)


def test_workspace_queries_do_not_need_a_gpu():
    need = _cabi.workspace_bytes(_cabi.OP_SEED_KNN, B=2, N=1000, C=128, S=100, k=40)      # distances + seed operand images
    assert 2 * 100 * 1000 * 4 + 2 * 131072 <= need <= 2 * 100 * 1000 * 4 + 2 * 131072 + 1024
    assert _cabi.workspace_bytes(_cabi.OP_PICK_SEEDS, B=1, N=5000) >= 5000 * 4
    assert _cabi.workspace_bytes(_cabi.OP_PICK_SEEDS, B=1, N=50000) >= 65536 * 8


def test_state_dict_layout_matches_reference():
    m = PointDSC(num_layers=12)
    sd = synth.make_state_dict(1234, num_layers=12)
    assert len(m.state_dict()) == 358
    assert sum(p.numel() for p in m.parameters()) == 1053667
    assert set(m.state_dict()) == set(sd)
    m.load_state_dict(sd, strict=True)
    for k, v in m.state_dict().items():
        assert v.shape == sd[k].shape, k


def test_cpu_tensors_are_refused():
    m = PointDSC(num_layers=1).eval()
    p = synth.make_pair(0, 64)
    with pytest.raises(RuntimeError, match="CUDA"):
        m({"corr_pos": p["corr_pos"], "src_keypts": p["src_keypts"], "tgt_keypts": p["tgt_keypts"], "testing": True})
    with pytest.raises(NotImplementedError):
        m({"corr_pos": p["corr_pos"], "src_keypts": p["src_keypts"], "tgt_keypts": p["tgt_keypts"]})


@pytest.mark.parametrize("case", ["g0_3dmatch_n256_l2", "g2_kitti_n800_l6"])
def test_host_mlps_match_oracle(case, monkeypatch):
    meta, g = load_golden(case)
    fam = synth.CONFIGS[meta["family"]]
    sd = synth.state_dict_for(meta, fam["sigma_d"])
    m = PointDSC(**meta["ctor"]).eval()
    m.load_state_dict(sd)

    def oracle_attn(q, k, v, src, tgt, sigma_spat, mode=0, out=None, ws=None):
        outs = []
        for b in range(q.shape[0]):
            _, compat = restate.spatial_compatibility(src[b:b + 1], tgt[b:b + 1], torch.tensor([sigma_spat]))
            msg = restate.nonlocal_message(q[b:b + 1].permute(0, 2, 1), k[b:b + 1].permute(0, 2, 1),
                                           v[b:b + 1].permute(0, 2, 1), compat)
            outs.append(msg.permute(0, 2, 1))
        return torch.cat(outs).contiguous()

    monkeypatch.setattr(_cabi, "nonlocal_attn", oracle_attn)
    with torch.no_grad():
        feat = m.encode(g["corr_pos"], g["src_keypts"], g["tgt_keypts"])
        conf = m.classify(feat)
    normed = torch.nn.functional.normalize(feat, p=2, dim=-1)
    assert (normed - g["normed_features"]).abs().max() < 2e-5
    scale = max(1.0, g["confidence"].abs().max().item())
    assert (conf - g["confidence"]).abs().max() < 1e-4 * scale


def test_argument_errors_come_back_as_codes_without_a_gpu():
    """Bad arguments are rejected before anything is launched: return code 1 and a message (SURVEY 8b error conventions)."""
    lib = _cabi.load()
    assert lib.pdsc_hybrid_search(None, 10, 0.1, 30, None, None, None, None, None, 0, None) == 1
    assert b"hybrid_search" in lib.pdsc_last_error()
    assert lib.pdsc_hybrid_search(1, 10, 0.1, 200, 1, None, None, None, 1, 0, None) == 1          # max_nn > 128
    assert lib.pdsc_voxel_down_sample(None, None, 5, 0.05, None, None, None, None, 0, None) == 1
    assert lib.pdsc_fpfh(None, None, 5, None, None, None, 100, None, None, None) == 1
    assert lib.pdsc_nn_match_f64(1, 1, 5, 5, 100, 1, None, 1, 0, None) == 1                        # D > 64
    assert lib.pdsc_icp_refine(None, None, None, 1, 10, 0.1, 30, 1e-6, 1e-6, None, None, None, None, 0, None) == 1
    assert lib.pdsc_icp_refine(1, 1, 1, 1, 10, -0.1, 30, 1e-6, 1e-6, None, None, None, 1, 0, None) == 1
    assert b"icp_refine" in lib.pdsc_last_error()
    assert lib.pdsc_pick_seeds(None, None, 1, 10, 0.1, 1, None, None, None, None, 0, None) == 1
    assert lib.pdsc_feature_head(None, 10, 128, 32, 32, None, None, None, None, None, None, 1e-12, None, None, None) == 1
    assert b"feature_head" in lib.pdsc_last_error()
    assert lib.pdsc_mlp_chain_direct(None, 10, 4, None, None, None, None, None, 2, None, None) == 1
    assert b"mlp_chain_direct" in lib.pdsc_last_error()


def test_mlp_chain_direct_refuses_bad_chains_before_touching_the_device():
    """Shape errors of pdsc_mlp_chain_direct are caught on the host (dims and relu are host arrays; the device pointers are
    never dereferenced on these paths, so dummy aligned addresses do)."""
    import ctypes
    lib = _cabi.load()
    I32 = ctypes.c_int32
    def call(n_layers, dims, residual, residual_layer):
        d = (I32 * len(dims))(*dims)
        r = (I32 * max(1, n_layers))(*([1] * max(1, n_layers)))
        return lib.pdsc_mlp_chain_direct(256, 10, n_layers, ctypes.cast(d, ctypes.c_void_p), ctypes.cast(r, ctypes.c_void_p), 512, 768,
                                         residual, residual_layer, 1024, None)
    assert call(5, [128] * 6, None, 0) == 1 and b"layers" in lib.pdsc_last_error()
    assert call(0, [128], None, 0) == 1
    assert call(2, [128, 96, 128], None, 0) == 1 and b"64 or 128" in lib.pdsc_last_error()
    assert call(2, [128, 64, 128], 1280, 2) == 1 and b"residual layer" in lib.pdsc_last_error()
    assert call(2, [128, 64, 128], 1281, 0) == 1 and b"aligned" in lib.pdsc_last_error()


def test_front_end_and_icp_refuse_cpu_tensors():
    from pointdsc_b200 import benchmark_utils, frontend
    pts = torch.rand(50, 3, dtype=torch.float64)
    with pytest.raises(RuntimeError, match="CUDA"):
        frontend.estimate_normal(pts, 0.1, 30)
    with pytest.raises(RuntimeError, match="CUDA"):
        benchmark_utils.icp_refine(torch.rand(1, 20, 3), torch.rand(1, 20, 3), torch.eye(4)[None])


def test_new_workspace_queries():
    n = 258342
    assert _cabi.workspace_bytes(_cabi.OP_HYBRID_SEARCH, N=n) >= n * 32
    assert _cabi.workspace_bytes(_cabi.OP_VOXEL_DOWN_SAMPLE, N=n, S=1 << 20) >= (1 << 20) * 16 + n * 8
    assert _cabi.workspace_bytes(_cabi.OP_NN_MATCH_F64, N=1000) >= 8 * 1000 * 16
    assert _cabi.workspace_bytes(_cabi.OP_ICP_REFINE, B=16, N=5000) >= 16 * 5000 * 28
