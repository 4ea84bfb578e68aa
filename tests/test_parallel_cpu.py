"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: pair sharding and the row-sharded
power iteration with its all-gather.  The CUDA kernels are replaced by oracle-based stand-ins
here (tests only); tests/test_gpu_kernels.py checks the kernels themselves."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import restate, synth
from pointdsc_b200 import parallel


def test_shard_ranges_partition_everything():
    for n in (0, 1, 7, 16, 4096, 50000):
        for world in (1, 2, 3, 4, 8):
            spans = [parallel.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    assert parallel.shard_pairs(10, 1, 4) == [3, 4, 5]


def _cpu_matvec(src, tgt, v, r0, r1, sigma, converged_flag=None, out=None):
    if converged_flag is not None and int(converged_flag) != 0:
        return out
    _, compat = restate.spatial_compatibility(src[None], tgt[None], torch.tensor([sigma]))
    M = compat[0].clone()
    M.fill_diagonal_(0)
    out.copy_(M[r0:r1] @ v)
    return out


def _cpu_normalize(y, v, flag, iters):
    if int(flag) != 0:
        return v
    vn = y / (torch.norm(y) + 1e-6)
    if torch.allclose(vn, v):
        flag.fill_(1)
    v.copy_(vn)
    iters += 1
    return v


def _worker(rank, world, port, N, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cfg = synth.CONFIGS["kitti"]
        p = synth.make_pair(200, N, **cfg)
        v, iters = parallel.power_iteration_full_sharded(p["src_keypts"][0], p["tgt_keypts"][0], cfg["sigma_d"], 10,
                                                         ops=(_cpu_matvec, _cpu_normalize))
        ref, used = restate.power_iteration_full(p["src_keypts"], p["tgt_keypts"], torch.tensor([cfg["sigma_d"]]), 10)
        out[rank] = (float((v - ref[0]).abs().max()), int(iters), used)
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


@pytest.mark.parametrize("N", [301, 600])
def test_row_sharded_power_iteration_world2_gloo(N):
    world = 2
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(world, _free_port(), N, out), nprocs=world, join=True)
        assert len(out) == world
        for rank in range(world):
            err, iters, used = out[rank]
            assert err < 1e-5, (rank, err)
            assert iters == used
