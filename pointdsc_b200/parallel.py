"""Multi-GPU pieces of the hot path (one process per GPU, torch.distributed).

* Registration pairs are independent: ``shard_pairs`` gives each rank its slice, no data-path
  collective (the pattern of the reference's test_parallel.sh:18-24 / DistributedSampler).
* One very large pair (N >= 50k, BASELINE config 4): the implicit N x N compatibility matrix is
  row-sharded.  Every power iteration each rank computes y[r0:r1] = M[r0:r1,:] v with the
  ``pdsc_compat_matvec`` kernel, the shards are all-gathered (NCCL over NVLink on GPUs; 4N bytes
  in total), and every rank normalises redundantly with ``pdsc_power_normalize`` -- which also
  evaluates the reference's allclose stopping rule on the device, so no second collective and
  no host sync is needed (cal_leading_eigenvector, models/PointDSC.py:338-358).
"""
from __future__ import annotations

import torch


def shard_range(n: int, rank: int, world: int):
    """Contiguous, balanced [begin, end) of n items for `rank` (first n % world ranks get one more)."""
    base, extra = divmod(n, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def shard_pairs(n_pairs: int, rank: int, world: int):
    return list(range(*shard_range(n_pairs, rank, world)))


def _default_ops():
    from . import _cabi
    return _cabi.compat_matvec, _cabi.power_normalize


def power_iteration_full_sharded(src, tgt, sigma_spat, num_iterations, group=None, ops=None):
    """Leading eigenvector of the zero-diagonal compatibility matrix of ONE pair, rows sharded over
    the ranks of `group`.  src,tgt: [N,3] (replicated on every rank).  Returns (v [N], iterations
    used, as a 1-element int32 tensor).  `ops` = (matvec, normalize) is injectable so the host
    logic can be exercised on CPU with gloo in the tests; the product path uses the CUDA kernels."""
    import torch.distributed as dist
    matvec, normalize = ops or _default_ops()
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    N = src.shape[0]
    r0, r1 = shard_range(N, rank, world)
    v = torch.ones(N, dtype=torch.float32, device=src.device)
    flag = torch.zeros(1, dtype=torch.int32, device=src.device)
    iters = torch.zeros(1, dtype=torch.int32, device=src.device)
    # all_gather needs equal shard sizes: pad every shard to the largest one
    width = (N + world - 1) // world
    y_full = torch.empty(world * width, dtype=torch.float32, device=src.device)
    y_loc = torch.zeros(width, dtype=torch.float32, device=src.device)
    bounds = [shard_range(N, r, world) for r in range(world)]
    y = torch.empty(N, dtype=torch.float32, device=src.device)
    for _ in range(num_iterations):
        matvec(src, tgt, v, r0, r1, sigma_spat, converged_flag=flag, out=y_loc[: r1 - r0])
        if world > 1:
            dist.all_gather_into_tensor(y_full, y_loc, group=group)
            if world * width == N:                      # equal shards: the gathered buffer is y
                y = y_full
            else:
                for r, (a, b) in enumerate(bounds):
                    y[a:b] = y_full[r * width: r * width + (b - a)]
        else:
            y[:] = y_loc[:N]
        normalize(y, v, flag, iters)
    return v, iters
