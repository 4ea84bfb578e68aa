"""pdsc_power_iter_full_static (no counterpart in the reference): the
no-allocation variant of pdsc_power_iter_full, whose scratch is a static device array inside the library.  Collected last
on purpose (file name), so that a failure here cannot hide another test under `pytest -x`.  Passes alone in a fresh process
and after the other tests: it depends on no state but the library's own static scratch, which every call re-initialises."""
import pytest
import torch

from oracle import synth


@pytest.mark.gpu
def test_static_scratch_power_iteration_matches_workspace_entry_point_twice():
    """Through the C ABI, two calls in a row (different sizes, so the second call finds the first one's contents in the
    shared scratch) against the existing entry point: same kernels, so the eigenvector must agree to the tolerance the
    power-iteration tests use (1e-5; in fact bit for bit) and the iteration counts exactly."""
    from pointdsc_b200 import _cabi
    for call, (N, fam, seed) in enumerate([(3000, "kitti", 31), (1777, "3dmatch", 32)]):
        cfg = synth.CONFIGS[fam]
        p = synth.make_pair(seed, N, **cfg)
        s, t = p["src_keypts"][0].cuda().contiguous(), p["tgt_keypts"][0].cuda().contiguous()
        ref_v, ref_it = _cabi.power_iter_full(s, t, cfg["sigma_d"], 10)
        v = torch.full((N,), float("nan"), device="cuda")            # outputs are the caller's
        it = torch.full((1,), -1, dtype=torch.int32, device="cuda")
        _cabi.power_iter_full_static(s, t, cfg["sigma_d"], 10, v, it)
        torch.cuda.synchronize()
        assert int(it[0]) == int(ref_it[0]), f"call {call}: iterations {int(it[0])} vs {int(ref_it[0])}"
        err = float((v - ref_v).abs().max())
        assert err <= 1e-5, f"call {call}: max |dv| = {err:.3e}"
        assert torch.isfinite(v).all()


@pytest.mark.gpu
def test_static_scratch_entry_point_makes_no_allocation_visible_to_torch():
    """torch's allocator statistics do not move across the call (the entry point itself calls no cudaMalloc by
    construction: see csrc/power.cu; this checks the binding adds none either)."""
    from pointdsc_b200 import _cabi
    cfg = synth.CONFIGS["3dmatch"]
    p = synth.make_pair(33, 1200, **cfg)
    s, t = p["src_keypts"][0].cuda().contiguous(), p["tgt_keypts"][0].cuda().contiguous()
    v = torch.empty(1200, device="cuda")
    it = torch.empty(1, dtype=torch.int32, device="cuda")
    _cabi.power_iter_full_static(s, t, cfg["sigma_d"], 10, v, it)     # first call: symbol address lookup
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats()["allocation.all.allocated"]
    _cabi.power_iter_full_static(s, t, cfg["sigma_d"], 10, v, it)
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats()["allocation.all.allocated"] == before


def test_static_scratch_entry_point_rejects_sizes_beyond_its_maximum():
    """Argument check only, so it runs without a GPU: N above PDSC_POWER_ITER_STATIC_MAX_N is refused with a clear error."""
    from pointdsc_b200 import _cabi
    lib = _cabi.load()
    rc = lib.pdsc_power_iter_full_static(1, 1, 1, _cabi.POWER_ITER_STATIC_MAX_N + 1, 0.1, 10, None, None)
    assert rc == 1
    msg = lib.pdsc_last_error().decode()
    assert "PDSC_POWER_ITER_STATIC_MAX_N" in msg and str(_cabi.POWER_ITER_STATIC_MAX_N + 1) in msg
    header = open(__import__("os").path.join(__import__("os").path.dirname(__file__), "..", "include", "pointdsc_b200.h")).read()
    assert f"#define PDSC_POWER_ITER_STATIC_MAX_N {_cabi.POWER_ITER_STATIC_MAX_N}" in header
