"""Row-sharded full-matrix power iteration over the GPUs of one box (BASELINE configs[3]; reference
models/PointDSC.py:338-358 on the N x N compatibility matrix, baseline_scripts/baseline_3DMatch.py:19-46): the NVLink P2P
transport (pdsc_compat_matvec_p2p + pdsc_power_normalize_p2p, parallel.power_iteration_full_p2p) must give the bits of the
single-GPU driver and of the NCCL all-gather path.  One process per GPU, so the test launches scripts/shard_power.py under
torchrun; skipped on a box with one GPU (tests/test_parallel_cpu.py covers the host logic with gloo there)."""
import json, os, subprocess, sys
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _run(world, n, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "scripts", "shard_power.py"), str(n)]
    out = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert lines, out.stdout[-2000:]
    return json.loads(lines[-1])


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on one box")
@pytest.mark.parametrize("n", [4096, 5003, 20000])      # a multiple of every tile, a ragged N, a several-wave grid
def test_p2p_power_iteration_is_bit_identical(n):
    world = min(torch.cuda.device_count(), 8)
    leg = _run(world, n, 29530 + n % 50)
    assert leg["n_gpus"] == world and leg["N"] == n
    assert "p2p_unavailable" not in leg, leg.get("p2p_unavailable")
    assert leg["transport"].startswith("nvlink p2p")
    assert leg["allgather_equals_p2p"] is True
    assert leg["bit_identical_to_1gpu"] is True
    assert 1 <= leg["iterations"] <= 10
